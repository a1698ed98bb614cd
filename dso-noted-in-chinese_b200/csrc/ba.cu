// K3-K5 placeholder translation unit (filled in by the BA kernels).
#include "common.cuh"
struct BAWorkspace { int dummy; };
int ba_create(dso_b200_ctx* ctx) { ctx->ba = nullptr; return 0; }
void ba_destroy(dso_b200_ctx* ctx) { (void)ctx; }
#define NOT_YET { dso_set_error("%s: BA kernels not built into this library yet", __func__); return DSO_B200_E_STATE; }
extern "C" int dso_b200_ba_set_window(dso_b200_ctx*, int, const int*, int, int, float, float, float, float, const float*) NOT_YET
extern "C" int dso_b200_ba_set_glue(dso_b200_ctx*, const dso_b200_precalc*, const double*, const double*, const float*, const float*) NOT_YET
extern "C" int dso_b200_ba_set_points(dso_b200_ctx*, int, const float*, const float*, const float*, const float*, const float*, const float*, const int*, const float*, const float*) NOT_YET
extern "C" int dso_b200_ba_set_residuals(dso_b200_ctx*, int, const int*, const int*, const int*, const double*, const unsigned char*, const unsigned char*) NOT_YET
extern "C" int dso_b200_ba_linearize(dso_b200_ctx*, int*, float*, float*, float*, float*, float*, double*) NOT_YET
extern "C" int dso_b200_ba_apply_res(dso_b200_ctx*, unsigned char*, float*) NOT_YET
extern "C" int dso_b200_ba_accumulate_A(dso_b200_ctx*, double*, double*, int*, float*, float*, float*) NOT_YET
extern "C" int dso_b200_ba_accumulate_SC(dso_b200_ctx*, double*, double*, float*, float*, float*) NOT_YET
extern "C" int dso_b200_ba_step(dso_b200_ctx*, double*, double*, double*, double*, double*, int*) NOT_YET
