// TEST INFRASTRUCTURE ONLY: see ../accumulators.hpp
#pragma once
#include "../accumulators.hpp"
