// coeffs_test.cpp -- prints the stage coefficients TimeStepper::setTimeStep (omega_b200/host/OmegaB200Stepper.h) forms
// for a time step and stepper type, 17 significant digits; tests/test_host_coeffs.py compares them bit for bit with
// capi.rk_coeffs, which the oracle-vs-reference tests pin to the reference's rational time arithmetic
// (src/infra/TimeMgr.cpp:193-283,382-391,747-766).  Host code only.
#include <cstdio>
#include <cstdlib>

#include "OmegaB200.h"
#include "OmegaB200Stepper.h"

using namespace OMEGA;

int main(int argc, char **argv) {
   if (argc < 3) return 2;
   const int Kind = std::atoi(argv[1]); // 0 RK4, 1 RK2, 2 FB (OmegaStepperType)
   const double Dt = std::strtod(argv[2], nullptr);
   TimeStepper S(Kind == 0 ? TimeStepperType::RungeKutta4 : Kind == 1 ? TimeStepperType::RungeKutta2 : TimeStepperType::ForwardBackward, Dt,
                 nullptr, nullptr, nullptr);
   std::printf("%d %.17g\n", S.Coeffs.NStages, S.Coeffs.Dt);
   for (int s = 0; s < S.Coeffs.NStages; ++s) std::printf("%.17g %.17g\n", S.Coeffs.A[s], S.Coeffs.B[s]);
   return 0;
}
