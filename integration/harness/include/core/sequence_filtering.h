#include "core/processing.h"
