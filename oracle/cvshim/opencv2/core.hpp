// TEST INFRASTRUCTURE: forwards to the OpenCV stand-in (see cvshim.hpp); not OpenCV code.
#include "cvshim.hpp"
