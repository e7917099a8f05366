// exact mode (R = double, Moeller-Trumbore in double): explicit instantiation
#include "gdpt_wave.cuh"
template int gdpt::run_wave<double>(gdpt_ctx *, gdpt::DScene<double>, gdpt::WaveParams);
