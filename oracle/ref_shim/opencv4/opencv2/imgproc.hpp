#include "shim.hpp"
