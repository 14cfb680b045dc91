// Stand-in for <pcl/search/impl/search.hpp>: included by NDT_H:45, nothing from it is used.  TEST INFRASTRUCTURE ONLY.
#pragma once
