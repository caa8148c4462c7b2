// ES3C (spike-and-slab) evaluator -- see es3c.cu
#pragma once
#include "common.cuh"
#include "estep.cuh"
namespace evo {
int es3c_launch_estep(const EstepDev &dv, const evo_estep_args *x, cudaStream_t st);
}
