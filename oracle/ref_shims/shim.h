// Force-included (-include) ahead of every reference TU compiled host-side.
// The reference relies on MSVC's transitive includes and on std::powf (absent from libstdc++ 13).
#pragma once
#include <cstdint>
#include <climits>
#include <cmath>
#include <cstdio>
#include <iostream>
namespace std { using ::powf; }
