#pragma once
#include "sim_world.cuh"
namespace ggl {
GGL_HD void collide_car_car(ArenaDev& A, int i, int j) { (void)A; (void)i; (void)j; }
}
