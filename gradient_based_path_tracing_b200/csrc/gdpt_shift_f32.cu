// shift-mapped gradpath (gdpt_shift.cuh), fast mode: explicit instantiation
#include "gdpt_shift.cuh"
template int gdpt::run_shift_wave<float>(gdpt_ctx *, gdpt::DScene<float>, gdpt::WaveParams);
