// Stand-in for tf/message_filter.h.
#pragma once
#include "tf.h"
