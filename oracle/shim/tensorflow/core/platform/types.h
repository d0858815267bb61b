// TEST INFRASTRUCTURE ONLY (oracle build).  Integer/string aliases the
// reference range coder pulls from TensorFlow (range_coder.h:23).
#ifndef ORACLE_SHIM_TF_TYPES_H_
#define ORACLE_SHIM_TF_TYPES_H_
#include <cstdint>
#include <limits>
#include <string>
namespace tensorflow {
typedef std::int8_t int8;
typedef std::int16_t int16;
typedef std::int32_t int32;
typedef std::int64_t int64;
typedef std::uint8_t uint8;
typedef std::uint16_t uint16;
typedef std::uint32_t uint32;
typedef std::uint64_t uint64;
using std::string;
typedef std::string tstring;
}  // namespace tensorflow
#endif
