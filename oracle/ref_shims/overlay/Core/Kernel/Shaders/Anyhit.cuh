// Case-fix forwarding header (build-side only; the reference tree stays read-only).
// BVHTraversal.cuh:4 includes "Core/Kernel/Shaders/Anyhit.cuh"; the file on disk is AnyHit.cuh.
#include "Core/Kernel/Shaders/AnyHit.cuh"
