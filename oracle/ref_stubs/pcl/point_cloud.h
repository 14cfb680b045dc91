// Stand-in for <pcl/point_cloud.h>.  TEST INFRASTRUCTURE ONLY (see pcl/point_types.h).
#pragma once
#include <pcl/point_types.h>
