// fast mode (R = float): explicit instantiation
#include "gdpt_wave.cuh"
template int gdpt::run_wave<float>(gdpt_ctx *, gdpt::DScene<float>, gdpt::WaveParams);
