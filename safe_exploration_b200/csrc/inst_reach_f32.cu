#define SE_T float
#include "inst_reach.cuh"
