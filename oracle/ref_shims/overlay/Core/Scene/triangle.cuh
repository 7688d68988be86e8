// Case-fix forwarding header: AnyHit.cuh:2 includes "Core/Scene/triangle.cuh"; the file is Triangle.cuh.
#include "Core/Scene/Triangle.cuh"
