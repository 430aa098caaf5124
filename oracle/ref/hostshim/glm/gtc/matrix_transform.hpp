// ORACLE - test infrastructure.  Part of the glm stand-in: everything is in glm/glm.hpp.
#pragma once
#include "../glm.hpp"
