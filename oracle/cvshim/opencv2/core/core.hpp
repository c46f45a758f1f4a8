// TEST INFRASTRUCTURE ONLY (oracle/_ref build): forwards to the OpenCV API subset in cvshim.h.
#include "cvshim.h"
