// ORACLE -- TEST INFRASTRUCTURE ONLY.  Stand-in for <opencv/cv.h>; see ../cvshim.hpp.
#include "../cvshim.hpp"
