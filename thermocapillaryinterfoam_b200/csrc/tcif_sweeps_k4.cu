// fixed-K sweeps for K = 4 (see tcif_sweeps_tu.inc)
#define TCIF_SWEEP_K 4
#include "tcif_sweeps_tu.inc"
