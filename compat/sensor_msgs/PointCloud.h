#pragma once
#include "PointCloud2.h"
