// forwards to the single stand-in header (oracle/ref_shim/pcl_standin.hpp)
#pragma once
#include "pcl_standin.hpp"
