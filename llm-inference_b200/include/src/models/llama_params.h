#pragma once
#include "src/models/llama/llama_params.h"  // the reference keeps an identical copy at this path
