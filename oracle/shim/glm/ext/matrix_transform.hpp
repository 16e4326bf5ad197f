// oracle shim: see ../glm.hpp
#include "../glm.hpp"
