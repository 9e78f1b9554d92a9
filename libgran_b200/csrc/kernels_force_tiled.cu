// kernels_force_tiled.cu -- placeholder: the tiled pair-parallel kernel is added in a later commit;
// until then variant 1 routes to the simple kernel.
#include "system.h"
namespace lg {
void launch_force(const ForceArgs& args, int mask, int variant, cudaStream_t st);
void launch_force_tiled(const ForceArgs& args, int mask, cudaStream_t st) { launch_force(args, mask, 0, st); }
}
