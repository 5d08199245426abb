// shim: see glm/shim_core.hpp
#pragma once
#include "../shim_core.hpp"
