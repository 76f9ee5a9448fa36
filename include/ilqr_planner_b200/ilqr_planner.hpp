// ilqr_planner.hpp — umbrella header of the C++ host API: namespace ilqr_planner::{sim, sys, solver} with the class names,
// constructors and solve() signatures of idiap/ilqr_planner, on top of the C ABI in include/ilqrb.h (link with -lilqrb).
#pragma once
#include "dense.hpp"
#include "keypoint.hpp"
#include "lqt.hpp"
#include "sd.hpp"
#include "serialize.hpp"
#include "sim.hpp"
#include "solver.hpp"
#include "system.hpp"
