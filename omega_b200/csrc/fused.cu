// fused.cu -- placeholder until the fused per-stage pipeline lands
#include "omega_dev.h"
#include "stepper.h"
namespace omega {
int fused_rk4_phase(OmegaStepper *, OmegaStateArrays *, const OmegaStepCoeffs *, int, int, cudaStream_t) {
   set_error("fused RK4 pipeline not built");
   return 3;
}
} // namespace omega
