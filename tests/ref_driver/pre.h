// TEST INFRASTRUCTURE, force-included (-include) in front of every reference source of tests/ref_driver/Makefile.
// examples/shared/Config.h hard-codes USE_CERES 1 (Ceres is not installed, and CeresSolverBase.h includes it by an
// absolute path of the author's machine); Config.h carries `#pragma once`, so including it here first and overriding the
// macro afterwards switches the Ceres code off without touching the reference's files.
#include "Config.h"
#undef USE_CERES
#define USE_CERES 0
