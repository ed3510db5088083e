// stand-in: graph.hpp includes this header but uses nothing from it in the functions compiled here
#pragma once
