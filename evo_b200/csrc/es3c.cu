#include "es3c.cuh"
namespace evo {
int es3c_launch_estep(const EstepDev &dv, const evo_estep_args *x, cudaStream_t st) {
  (void)dv; (void)x; (void)st;
  set_error("ES3C evaluator not built yet");
  return EVO_ERR_UNSUPPORTED;
}
}
