// Stand-in: the reference's Scancontext.h includes this header and uses nothing from it.  TEST INFRASTRUCTURE ONLY (see pcl/point_types.h).
#pragma once
