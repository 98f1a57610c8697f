// TEMPORARY stubs until conv_gemm.cu lands.
#include "hpe_common.cuh"
using namespace hpe;
extern "C" int hpe_conv_plan_create(const hpe_conv_desc*, hpe_conv_plan**) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
extern "C" void hpe_conv_plan_destroy(hpe_conv_plan*) {}
extern "C" size_t hpe_conv_packed_weight_elems(const hpe_conv_desc*) { return 0; }
extern "C" int hpe_conv_pack_weight(const hpe_conv_desc*, const float*, void*, void*) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
extern "C" int hpe_conv_run(hpe_conv_plan*, const void*, const void*, const float*, const float*, const void*, void*, void*) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
extern "C" int hpe_conv_out_hw(const hpe_conv_desc*, int*, int*) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
extern "C" size_t hpe_preprocess_out_elems(int, int, int) { return 0; }
extern "C" int hpe_preprocess(const float*, int, int, int, void*, void*) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
extern "C" int hpe_maxpool3x3s2(const void*, int, int, int, int, void*, void*) { return fail(HPE_EUNSUPPORTED, "conv not built yet"); }
