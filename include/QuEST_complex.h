/* quest-b200: the user-side complex-number convenience header of the QuEST distribution
 * (reference: QuEST/src/QuEST_complex.h:19-86, shipped next to QuEST.h and included by user programs only - no
 * library source uses it).  It gives user code one native complex type, `qcomp`, whose width follows QuEST_PREC, and
 * the two conversions to and from the `Complex` struct the API takes.  The names below are the interface user programs
 * are written against; nothing here touches the GPU.
 *
 *   C   : qcomp is `float|double|long double complex` (<complex.h>), built with qcomp(re, im)
 *   C++ : qcomp is std::complex<float|double|long double>,          built with qcomp(re, im)
 *   both: creal/cimag/carg/cabs work on a qcomp; toComplex(z) -> Complex, fromComplex(c) -> qcomp
 */
#ifndef QUEST_COMPLEX_H
#define QUEST_COMPLEX_H

#include "QuEST_precision.h"

#ifdef __cplusplus

#include <cmath>
#include <complex>

/* the reference header opens namespace std for its users; programs written against it rely on that */
using namespace std;

typedef complex<float> float_complex;
typedef complex<double> double_complex;
typedef complex<long double> long_double_complex;

/* C99 spellings of the accessors, so that one source text serves both languages */
#define creal(z) real(z)
#define cimag(z) imag(z)
#define carg(z) arg(z)
#define cabs(z) abs(z)

#else /* C99 */

#include <tgmath.h> /* pulls in <math.h> and <complex.h>, with type-generic maths */

typedef float complex float_complex;
typedef double complex double_complex;
typedef long double complex long_double_complex;

/* constructor-style spellings matching the C++ side: double_complex(1, 2) etc. */
#define float_complex(re, im) ((float)(re) + ((float)(im)) * I)
#define double_complex(re, im) ((double)(re) + ((double)(im)) * I)
#define long_double_complex(re, im) ((long double)(re) + ((long double)(im)) * I)

#endif

/* one precision-agnostic name, the complex analogue of qreal */
#if QuEST_PREC == 1
#define qcomp float_complex
#elif QuEST_PREC == 2
#define qcomp double_complex
#elif QuEST_PREC == 4
#define qcomp long_double_complex
#endif

/* qcomp <-> the Complex struct of QuEST.h */
#ifdef __cplusplus
#define toComplex(z) (Complex{(qreal)creal(z), (qreal)cimag(z)})
#else
#define toComplex(z) ((Complex){.real = creal(z), .imag = cimag(z)})
#endif
#define fromComplex(c) qcomp((c).real, (c).imag)

#endif /* QUEST_COMPLEX_H */
