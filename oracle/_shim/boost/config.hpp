// TEST INFRASTRUCTURE ONLY: see boost/graph/adjacency_list.hpp in this directory
