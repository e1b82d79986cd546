// OBJLoader.hpp -- forwarding header: keeps `#include "OBJLoader.hpp"` of code written against AFIDclan/cuda-raytracing compiling.
// The whole drop-in surface lives in raytracer.hpp.
#pragma once
#include "raytracer.hpp"
