// Process-side plumbing of the class shims: one librgbdfe context per calling thread
// (the reference runs the hot path from two threads: tracking, System/Tracking.cpp:39, and the
// pose graph, Solver/PoseGraph.cpp:56), created lazily on first use.  No CPU fallback: when the
// context cannot be created (no CUDA device, library missing) the process stops with a message,
// as the reference's own fatal paths do (std::terminate, Features/Extractor.cpp:39).
#ifndef RGBDFE_HOST_H
#define RGBDFE_HOST_H

#include "../../include/rgbdfe.h"

namespace rgbdfe_host {

// Capacity of the lazily created contexts; call before the first shim call to change it.
// Defaults: 1920 x 1080 images, 4000 features (BASELINE configs[2] is the largest case).
void set_capacity(int max_w, int max_h, int max_feats);
void set_device(int device);

// The calling thread's context.
rgbdfe_ctx* ctx();

// Which extractor object last configured the calling thread's context (ORB parameters are context state).
const void*& config_owner();

// Device-resident copy of a Frame's descriptors (n x 32 B) and mvKeys3Dc (n x 12 B) in the calling thread's context,
// uploaded on first use.  Matcher::match(ref, cur) and RansacSE3::compute(ref, cur, ...) name the same two Frames and
// `cur` is the next call's `ref`, so each set goes up once instead of three times.  The key is the CONTENT (a 64-bit
// hash of both arrays and n), never an address: a Frame freed and reallocated at the same address cannot alias.
// A small LRU (8 frames); returns nullptr on error (reported).
rgbdfe_frame* resident(const void* desc, const void* xyz, int n);

// Report `rc` (an RGBDFE_* code) of a failed call on stderr; returns rc.
int report(const char* where, int rc);

// Random's stream lives in the context of the thread that draws from it.  Random::initSeed()
// seeds it from time(NULL) once per process (System/Random.cpp:7-13); tests pin it here.
void seed(unsigned s);
bool seeded();

} // namespace rgbdfe_host

#endif
