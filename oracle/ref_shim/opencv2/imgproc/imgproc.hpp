// ORACLE -- TEST INFRASTRUCTURE ONLY.  Stand-in for the OpenCV header of this name; see ../../cvshim.hpp.
#include "../../cvshim.hpp"
