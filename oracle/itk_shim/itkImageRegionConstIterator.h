#include "itk_shim.h"
