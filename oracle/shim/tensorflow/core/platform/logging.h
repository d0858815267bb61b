// TEST INFRASTRUCTURE ONLY (oracle build).  CHECK/DCHECK macros used by the
// reference range coder (range_coder.cc:28).  CHECKs abort with a message;
// DCHECKs are active unless NDEBUG (TF's opt build compiles them out).
#ifndef ORACLE_SHIM_TF_LOGGING_H_
#define ORACLE_SHIM_TF_LOGGING_H_
#include <cstdio>
#include <cstdlib>
#define ORACLE_CHECK_IMPL(cond, txt)                                        \
  do {                                                                      \
    if (!(cond)) {                                                          \
      std::fprintf(stderr, "%s:%d CHECK failed: %s\n", __FILE__, __LINE__,  \
                   txt);                                                    \
      std::abort();                                                         \
    }                                                                       \
  } while (0)
#define CHECK(c) ORACLE_CHECK_IMPL((c), #c)
#define CHECK_EQ(a, b) ORACLE_CHECK_IMPL((a) == (b), #a " == " #b)
#define CHECK_NE(a, b) ORACLE_CHECK_IMPL((a) != (b), #a " != " #b)
#define CHECK_LT(a, b) ORACLE_CHECK_IMPL((a) < (b), #a " < " #b)
#define CHECK_LE(a, b) ORACLE_CHECK_IMPL((a) <= (b), #a " <= " #b)
#define CHECK_GT(a, b) ORACLE_CHECK_IMPL((a) > (b), #a " > " #b)
#define CHECK_GE(a, b) ORACLE_CHECK_IMPL((a) >= (b), #a " >= " #b)
#ifdef NDEBUG
#define DCHECK(c) do { } while (0)
#define DCHECK_EQ(a, b) do { } while (0)
#define DCHECK_NE(a, b) do { } while (0)
#define DCHECK_LT(a, b) do { } while (0)
#define DCHECK_LE(a, b) do { } while (0)
#define DCHECK_GT(a, b) do { } while (0)
#define DCHECK_GE(a, b) do { } while (0)
#else
#define DCHECK(c) CHECK(c)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) CHECK_NE(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#define DCHECK_LE(a, b) CHECK_LE(a, b)
#define DCHECK_GT(a, b) CHECK_GT(a, b)
#define DCHECK_GE(a, b) CHECK_GE(a, b)
#endif
#endif
