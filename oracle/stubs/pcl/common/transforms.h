#pragma once
// real PCL pulls point_cloud.h in through this header
#include <pcl/point_cloud.h>
