#pragma once
#include "sim_tick.cuh"
