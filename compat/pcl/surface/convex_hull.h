// Stand-in: see concave_hull.h.
#pragma once
#include "concave_hull.h"
