// forwards to the single stand-in header (oracle/ref_shim/pcl_standin.hpp); the reference includes GICP but never uses it
#pragma once
#include "pcl_standin.hpp"
