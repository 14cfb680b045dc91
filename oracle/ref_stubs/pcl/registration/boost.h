// Stand-in for <pcl/registration/boost.h>.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <pcl/point_types.h>
