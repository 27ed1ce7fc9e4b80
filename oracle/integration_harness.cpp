// oracle/integration_harness.cpp -- TEST INFRASTRUCTURE. The one thing the compiled integration patch
// (oracle/build_integration.py) needs besides the reference's sources: the mcpt context the viewer would own
// (RTViewer::ctx_, created in init_ogl per INTEGRATION.md). The test hands it in before BVH_GPU_Scene::finalize() runs.
#include <mcpt.h>

extern "C" {
mcpt_ctx* mcpt_integration_ctx = nullptr;
void integration_set_ctx(mcpt_ctx* c) { mcpt_integration_ctx = c; }
}
