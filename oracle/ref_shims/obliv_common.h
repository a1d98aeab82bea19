/* TEST INFRASTRUCTURE: shim for the absent Obliv-C header of the same name. */
#pragma once
#include "p1shim.h"
