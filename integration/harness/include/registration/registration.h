/* TEST HARNESS ONLY — src/registration/registration.h:222 */
#ifndef HARNESS_REGISTRATION_H
#define HARNESS_REGISTRATION_H
#include "core/siril.h"
void translation_from_H(Homography H, double *dx, double *dy);
#endif
