/* cugraph_c/algorithms.h - umbrella header, as cpp/include/cugraph_c/algorithms.h */
#pragma once

#include <cugraph_c/centrality_algorithms.h>
