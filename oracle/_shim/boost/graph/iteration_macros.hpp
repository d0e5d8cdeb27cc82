// TEST INFRASTRUCTURE ONLY: see adjacency_list.hpp in this directory
#include "adjacency_list.hpp"
