// shift-mapped gradpath (gdpt_shift.cuh), exact mode: explicit instantiation
#include "gdpt_shift.cuh"
template int gdpt::run_shift_wave<double>(gdpt_ctx *, gdpt::DScene<double>, gdpt::WaveParams);
