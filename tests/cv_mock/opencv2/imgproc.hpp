// see core.hpp: the mock declares convexHull / contourArea there
#pragma once
#include "core.hpp"
