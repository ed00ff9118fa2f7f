/* drop-in for the reference header of the same name: see b200_ref_compat.h */
#pragma once
#include "b200_ref_compat.h"
