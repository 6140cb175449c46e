// Stand-in for boost/filesystem.hpp (unused by the compiled sources).
#pragma once
