// temporary: entry points implemented in frame.cu / stereo.cu
#include "matcher.cuh"
using namespace ydorb;
extern "C" {
int ydorb_frame_create(YdMatcher*, const YdFrameGeom*, const YdKeyPoint*, const uint8_t*, const float*, int, YdFrame**) { set_error("not implemented"); return YDORB_E_STATE; }
void ydorb_frame_destroy(YdFrame*) {}
int ydorb_frame_keypoints_in_area(YdFrame*, float, float, float, int, int, int32_t*, int, int32_t*) { set_error("not implemented"); return YDORB_E_STATE; }
int ydorb_search_by_projection(YdFrame*, const YdMapPointProj*, const uint8_t*, const uint8_t*, int, const uint8_t*, float, float, int32_t*, int32_t*) { set_error("not implemented"); return YDORB_E_STATE; }
int ydorb_stereo_match(YdMatcher*, const YdExtractor*, int, const YdExtractor*, int, const YdKeyPoint*, const uint8_t*, int, const YdKeyPoint*, const uint8_t*, int, float, float, int, float*, float*) { set_error("not implemented"); return YDORB_E_STATE; }
}
