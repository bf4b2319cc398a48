// forwards to the single stand-in header (oracle/ref_shim/tf2_standin.hpp)
#pragma once
#include "tf2_standin.hpp"
