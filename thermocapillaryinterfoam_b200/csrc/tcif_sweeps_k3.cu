// fixed-K sweeps for K = 3 (see tcif_sweeps_tu.inc)
#define TCIF_SWEEP_K 3
#include "tcif_sweeps_tu.inc"
