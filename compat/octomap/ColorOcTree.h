// Stand-in: see octomap.h.
#pragma once
#include "octomap.h"
