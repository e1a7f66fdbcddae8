#include "pbrt.h"
