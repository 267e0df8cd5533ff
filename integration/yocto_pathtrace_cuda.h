// yocto_pathtrace_cuda.h — the three calls the CUDA adapter adds next to the reference's
// yocto_pathtrace.h API (which it otherwise keeps unchanged; see yocto_pathtrace_cuda.cpp).
#ifndef YOCTO_PATHTRACE_CUDA_H_
#define YOCTO_PATHTRACE_CUDA_H_

#include <yocto_pathtrace/yocto_pathtrace.h>

#include <string>

namespace yocto {

// The scene file load_scene read. Needed only for scenes with analytic SDFs: sdf_data keeps a
// std::function closure, not {type, parameters} (yocto_sceneio.cpp:3685-3728), so the adapter
// re-reads the `sdfunctions` section and checks it against the closures.
void pathtrace_cuda_scene_file(const std::string& filename);

// Refreshes state.image / hits / rngs from the device (get_render alone needs no sync).
void pathtrace_cuda_sync(pathtrace_state& state);

// Frees the device mirrors of a pathtrace_state (call before the state is destroyed).
void pathtrace_cuda_release(const pathtrace_state& state);

// implementation detail of pathtrace_samples (the caller holds the adapter's lock)
void pathtrace_samples_unlocked(pathtrace_state& state, const scene_data& scene,
    const bvh_scene& bvh, const pathtrace_lights& lights, const pathtrace_params& params);

}  // namespace yocto

#endif
