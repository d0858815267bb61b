#include "tensorflow/core/framework/op_kernel.h"  // stub: see that file
