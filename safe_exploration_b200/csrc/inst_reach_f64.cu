#define SE_T double
#include "inst_reach.cuh"
