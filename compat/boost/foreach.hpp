// Stand-in for boost/foreach.hpp (unused by the compiled sources).
#pragma once
