// compat: glog/logging.h — VLOG is swallowed (see paddle/fluid/platform/enforce.h).
#pragma once
#include "paddle/fluid/platform/enforce.h"
