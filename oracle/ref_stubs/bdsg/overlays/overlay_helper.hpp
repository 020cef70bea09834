// libbdsg stand-in.  TEST INFRASTRUCTURE (see ../../README.md).
#pragma once
#include "../snarl_distance_index.hpp"
