// Stand-in: see Simple_cartesian.h.
#pragma once
#include "Simple_cartesian.h"
