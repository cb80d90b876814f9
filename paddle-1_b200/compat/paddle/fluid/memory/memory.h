#pragma once
#include "paddle/fluid/memory/memcpy.h"
