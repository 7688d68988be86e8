// Case-fix forwarding header: Bounds.cuh:3 includes "Hitpayload.cuh"; the file is Core/HitPayload.cuh.
#include "Core/HitPayload.cuh"
