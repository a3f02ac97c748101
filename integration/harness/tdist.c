/* TEST HARNESS ONLY — gsl_cdf_tdist_Pinv for the harness (GSL is absent here): the Student-t
 * distribution function through the regularized incomplete beta function (Lentz's continued
 * fraction), inverted by bisection.  tests/test_tdist_quantile.py pins the same construction on scipy. */
#include <math.h>

static double betacf(double a, double b, double x) {
	const double tiny = 1e-300;
	double qab = a + b, qap = a + 1.0, qam = a - 1.0, c = 1.0, d = 1.0 - qab * x / qap, h;
	if (fabs(d) < tiny) d = tiny;
	d = 1.0 / d; h = d;
	for (int m = 1; m < 400; m++) {
		const int m2 = 2 * m;
		double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
		d = 1.0 + aa * d; if (fabs(d) < tiny) d = tiny;
		c = 1.0 + aa / c; if (fabs(c) < tiny) c = tiny;
		d = 1.0 / d; h *= d * c;
		aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
		d = 1.0 + aa * d; if (fabs(d) < tiny) d = tiny;
		c = 1.0 + aa / c; if (fabs(c) < tiny) c = tiny;
		d = 1.0 / d;
		const double de = d * c;
		h *= de;
		if (fabs(de - 1.0) < 1e-16) break;
	}
	return h;
}
static double betai(double a, double b, double x) {
	if (x <= 0.0) return 0.0;
	if (x >= 1.0) return 1.0;
	const double bt = exp(lgamma(a + b) - lgamma(a) - lgamma(b) + a * log(x) + b * log1p(-x));
	if (x < (a + 1.0) / (a + b + 2.0)) return bt * betacf(a, b, x) / a;
	return 1.0 - bt * betacf(b, a, 1.0 - x) / b;
}
static double t_cdf(double t, double nu) {
	const double tail = 0.5 * betai(0.5 * nu, 0.5, nu / (nu + t * t));
	return t > 0 ? 1.0 - tail : tail;
}
double gsl_cdf_tdist_Pinv(double P, double nu) {
	if (P == 0.5) return 0.0;
	if (P < 0.5) return -gsl_cdf_tdist_Pinv(1.0 - P, nu);
	double lo = 0.0, hi = 1e6;
	for (int i = 0; i < 200; i++) {
		const double mid = 0.5 * (lo + hi);
		if (t_cdf(mid, nu) < P) lo = mid; else hi = mid;
		if (hi - lo <= 1e-15 * hi) break;
	}
	return 0.5 * (lo + hi);
}
