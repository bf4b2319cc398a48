// forwards to the single mock header (tests/mock_ros2/mock_ros2.hpp)
#pragma once
#include "mock_ros2.hpp"
