// Shim so that the reference's dependency-free header
//   /root/reference/src/continuous-validation/intervals.hh
// compiles in place without Eigen/Boost/pinocchio.  Only the two typedefs
// that header needs; values as in reference include/hpp/core/fwd.hh:174,181.
#pragma once
#include <sstream>
#include <string>
#include <utility>
namespace hpp { namespace core {
typedef double value_type;
typedef std::pair<value_type, value_type> interval_t;
}}
