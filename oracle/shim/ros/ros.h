// Shim for <ros/ros.h>: TEST INFRASTRUCTURE ONLY (oracle/_ref build).
// The reference hot path uses ROS solely for three logging macros
// (ccicp2d/src/icp.cpp:22,31,39,90,96,101; icpPointToPoint.cpp:22,129).
// They are compiled to no-ops that still type-check their stream argument.
#pragma once
#include <iostream>
#include <sstream>
#define CCICP_SHIM_LOG(x) do { if (false) { std::ostringstream _s; _s << x; } } while (0)
#define ROS_ERROR_STREAM(x) CCICP_SHIM_LOG(x)
#define ROS_WARN_STREAM(x)  CCICP_SHIM_LOG(x)
#define ROS_INFO_STREAM(x)  CCICP_SHIM_LOG(x)
