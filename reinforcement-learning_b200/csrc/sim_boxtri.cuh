#pragma once
#include "sim_world.cuh"
namespace ggl {
GGL_HD void collide_car_mesh(ArenaDev& A, const WorldDev& W, int ci) { (void)A; (void)W; (void)ci; }
}
