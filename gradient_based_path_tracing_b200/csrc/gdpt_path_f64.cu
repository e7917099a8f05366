// `path` integrator, exact mode (R = double, compiled without FMA contraction): explicit instantiation
#include "gdpt_path_wave.cuh"
template int gdpt::run_path_wave<double>(gdpt_ctx *, gdpt::DScene<double>, gdpt::WaveParams);
