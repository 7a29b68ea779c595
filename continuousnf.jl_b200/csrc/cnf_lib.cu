// cnf_lib.cu — single translation unit of libcnfb200.so (keeps __constant__ tables and
// templates in one module; no relocatable device code needed).
#include "cnf_solver.cu"
#include "cnf_stage_generic.cu"
#include "cnf_stage_narrow.cu"
#include "cnf_vector.cu"
#include "cnf_stage_tc.cu"
#include "cnf_backward_tc.cu"
#include "cnf_backward.cu"
#include "cnf_rng.cu"
#include "cnf_nccl.cu"
#include "cnf_api.cu"
#include "cnf_tile.cu"
#include "cnf_fit.cu"
