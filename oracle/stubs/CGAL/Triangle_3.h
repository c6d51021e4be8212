#pragma once
namespace CGAL { template <class K> struct Triangle_3 {}; }
