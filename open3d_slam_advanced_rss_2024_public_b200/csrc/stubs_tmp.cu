// TEMPORARY: entry points not implemented yet (replaced by filters.cu / map.cu)
#include "common.cuh"
extern "C" {
int32_t b200reg_surface_normals(b200reg_ctx* ctx, b200reg_cloud*, int32_t, float, float, const float*, float*, float*, float*, int32_t*) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
int32_t b200reg_map_create(b200reg_ctx* ctx, double, b200reg_map**) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
void b200reg_map_destroy(b200reg_ctx*, b200reg_map*) {}
int64_t b200reg_map_size(const b200reg_map*) { return 0; }
int32_t b200reg_map_insert(b200reg_ctx* ctx, b200reg_map*, const double*, const double*, int64_t, const double*, const b200reg_crop_spec*) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
int32_t b200reg_map_crop(b200reg_ctx* ctx, const b200reg_map*, const b200reg_crop_spec*, b200reg_cloud**) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
int32_t b200reg_map_download(b200reg_ctx* ctx, const b200reg_map*, double*, double*) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
int32_t b200reg_crop_f64(b200reg_ctx* ctx, const b200reg_crop_spec*, const double*, const double*, int64_t, double*, double*, int64_t*) { return setErr(ctx, B200REG_E_INVALID_PARAM, "not implemented"); }
}
