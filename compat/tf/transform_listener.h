// Stand-in for tf/transform_listener.h.
#pragma once
#include "tf.h"
