// `path` integrator, fast mode (R = float): explicit instantiation
#include "gdpt_path_wave.cuh"
template int gdpt::run_path_wave<float>(gdpt_ctx *, gdpt::DScene<float>, gdpt::WaveParams);
