// oracle-only forwarding header for the glm stand-in (see glm_shim.hpp)
#include "glm_shim.hpp"
