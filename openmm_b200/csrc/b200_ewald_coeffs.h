// Rational approximations produced by tools/fit_ewald_kernels.py (valid for t = (alpha r)^2 in [0,16]).
//   F(t) = (erf(z)/z - 2/sqrt(pi) exp(-t))/t   -> force:  erfc(ar)/r + 2a/sqrt(pi) exp(-a^2 r^2) = 1/r - a^3 r^2 F(t)
//   V(t) = erf(z)/z                            -> energy: erfc(ar)/r = 1/r - a V(t)
// They replace the exp + Abramowitz-Stegun erfc polynomial of the reference's direct-space snippet
// (platforms/common/src/kernels/coulombLennardJones.cc:9-23) with one division and no MUFU.EX2.
// max relative error: F (6,5) 6.2e-8 in exact arithmetic, 3.4e-7 evaluated in fp32; V (6,6) 9e-10 / 3.2e-7.
#pragma once
#define B200_EWALD_TMAX 16.0f
#define F_N0 7.522528250e-01f
#define F_N1 -3.588437124e-02f
#define F_N2 1.807796133e-02f
#define F_N3 -2.020375374e-04f
#define F_N4 5.036963867e-05f
#define F_N5 -1.028338385e-06f
#define F_N6 1.081107626e-08f
#define F_D1 5.522995154e-01f
#define F_D2 1.411143827e-01f
#define F_D3 2.163091483e-02f
#define F_D4 2.096388751e-03f
#define F_D5 1.302120108e-04f
#define V_N0 1.128379168e+00f
#define V_N1 1.669044217e-01f
#define V_N2 5.093868480e-02f
#define V_N3 2.956066784e-03f
#define V_N4 3.057635111e-04f
#define V_N5 -2.883694124e-07f
#define V_N6 -4.483770719e-08f
#define V_D1 4.812485646e-01f
#define V_D2 1.055591566e-01f
#define V_D3 1.349151546e-02f
#define V_D4 1.039880177e-03f
#define V_D5 4.100137982e-05f
#define V_D6 -6.200463236e-07f
