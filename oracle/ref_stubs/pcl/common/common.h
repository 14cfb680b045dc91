// Stand-in for <pcl/common/common.h>.  TEST INFRASTRUCTURE ONLY -- see pcl/point_types.h.
#pragma once
#include <pcl/point_types.h>
