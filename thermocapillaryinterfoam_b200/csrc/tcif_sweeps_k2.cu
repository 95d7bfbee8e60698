// fixed-K sweeps for K = 2 (see tcif_sweeps_tu.inc)
#define TCIF_SWEEP_K 2
#include "tcif_sweeps_tu.inc"
