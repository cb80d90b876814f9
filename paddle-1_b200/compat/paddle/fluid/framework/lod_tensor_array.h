#pragma once
#include <vector>
#include "paddle/fluid/framework/lod_tensor.h"
namespace paddle { namespace framework { using LoDTensorArray = std::vector<LoDTensor>; } }
