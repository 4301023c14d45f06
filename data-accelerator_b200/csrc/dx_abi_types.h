// dx_abi_types.h — pulls the public C ABI declarations into the implementation files.
#pragma once
#include "dxgpu.h"
