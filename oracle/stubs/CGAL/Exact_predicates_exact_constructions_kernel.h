#pragma once
namespace CGAL { struct Exact_predicates_exact_constructions_kernel {}; }
