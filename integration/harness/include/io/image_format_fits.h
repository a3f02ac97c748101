/* TEST HARNESS ONLY — src/io/image_format_fits.h:22-60 */
#ifndef HARNESS_FITS_H
#define HARNESS_FITS_H
#include "core/siril.h"
data_type get_data_type(int bitpix);
void clearfits(fits *);
int new_fit_image(fits **fit, int width, int height, int nblayer, data_type type);
#endif
