/* TEST HARNESS ONLY — gsl/gsl_cdf.h (GSL is a system dependency of Siril, absent here) */
#ifndef HARNESS_GSL_CDF_H
#define HARNESS_GSL_CDF_H
double gsl_cdf_tdist_Pinv(double P, double nu);
#endif
