#include "../shim.hpp"
