// OpenCV stand-in for building oracle/_ref (test infrastructure): everything lives in mini_cv.hpp.
#pragma once
#include "mini_cv.hpp"
