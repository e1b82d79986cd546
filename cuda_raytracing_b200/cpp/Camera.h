// Camera.h -- forwarding header: keeps `#include "Camera.h"` of code written against AFIDclan/cuda-raytracing compiling.
// The whole drop-in surface lives in raytracer.hpp.
#pragma once
#include "raytracer.hpp"
