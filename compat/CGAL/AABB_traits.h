// Stand-in: see AABB_tree.h.
#pragma once
#include "AABB_tree.h"
