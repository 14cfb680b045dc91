// Stand-in for <pcl/filters/boost.h> (boost shims live in pcl/point_types.h).  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <pcl/point_types.h>
