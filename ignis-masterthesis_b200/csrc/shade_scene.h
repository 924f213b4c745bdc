// shade_scene.h -- host-visible declaration of the shading view (the device functions live in shade_lib.cuh).
#pragma once
#include "shade_lib.cuh"
