/* detmath_interpose.c -- TEST INFRASTRUCTURE ONLY (see nbody_oracle.c header).
 *
 * Defines atan2, sin, cos and sincos with include/nbody_detmath.h so that the
 * UNMODIFIED reference source (/root/reference/nbody/nbody-seq.c, compiled where
 * it lies by oracle/Makefile) resolves its trig calls (nbody-seq.c:91,97,98,
 * 123-125) here instead of in libm.  sqrt stays the inlined IEEE sqrtsd.
 * Build with -fno-builtin -ffp-contract=off and link this object before -lm.
 */
#include "../include/nbody_detmath.h"

static const double dm_tab[NBODY_DM_ATAN_TAB_LEN] = NBODY_DM_ATAN_TAB_INIT;

double atan2(double y, double x) { return nbody_dm_atan2(y, x, dm_tab); }
double sin(double a) { double s, c; nbody_dm_sincos(a, &s, &c); return s; }
double cos(double a) { double s, c; nbody_dm_sincos(a, &s, &c); return c; }
void sincos(double a, double *s, double *c) { nbody_dm_sincos(a, s, c); }
