// MeshPrimitive.h -- forwarding header: keeps `#include "MeshPrimitive.h"` of code written against AFIDclan/cuda-raytracing compiling.
// The whole drop-in surface lives in raytracer.hpp.
#pragma once
#include "raytracer.hpp"
