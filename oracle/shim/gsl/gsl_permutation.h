/* TEST INFRASTRUCTURE ONLY — see gsl_linalg.h */
#include "gsl_linalg.h"
