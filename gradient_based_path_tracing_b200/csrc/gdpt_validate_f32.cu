// validation launchers, R = float: explicit instantiation
#include "gdpt_validate.cuh"
namespace gdpt {
template int dbg_sample_primary<float>(gdpt_ctx *, const DScene<float> &, int, const double *, double *);
template int dbg_intersect<float>(gdpt_ctx *, const DScene<float> &, int, const double *, const double *, int32_t *, double *);
template int dbg_bsdf<float>(gdpt_ctx *, const DScene<float> &, int, int, const double *, const double *, const double *, double *);
template int dbg_sample_light<float>(gdpt_ctx *, const DScene<float> &, int, const double *, const double *, double *);
template int dbg_aux<float>(gdpt_ctx *, const DScene<float> &, int, double *);
}
