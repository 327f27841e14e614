/* quest-b200: precision selection for the QuEST C API.
 *
 * Replaces (same macro names, same values) reference QuEST/include/QuEST_precision.h:27-61.
 * QuEST_PREC is a compile-time switch: 1 = float, 2 = double (default). The B200 engine
 * ships PREC 1 and 2 only; quad precision (4) has no GPU arithmetic and is rejected here,
 * as the reference's own GPU build rejects it (QuEST/CMakeLists.txt:66-70).
 */
#ifndef QUEST_PRECISION_H
#define QUEST_PRECISION_H

#include <math.h>

#ifndef QuEST_PREC
#define QuEST_PREC 2
#endif

#if QuEST_PREC == 1
#define qreal float
#define MPI_QuEST_REAL MPI_FLOAT
#define MPI_MAX_AMPS_IN_MSG (1LL << 29)
#define REAL_STRING_FORMAT "%.8f"
#define REAL_QASM_FORMAT "%.8g"
#define REAL_EPS 1e-5
#define absReal(X) fabs(X)
#elif QuEST_PREC == 2
#define qreal double
#define MPI_QuEST_REAL MPI_DOUBLE
#define MPI_MAX_AMPS_IN_MSG (1LL << 28)
#define REAL_STRING_FORMAT "%.14f"
#define REAL_QASM_FORMAT "%.14g"
#define REAL_EPS 1e-13
#define absReal(X) fabs(X)
#else
#error "quest-b200 supports QuEST_PREC 1 (float) and 2 (double) only"
#endif

#endif /* QUEST_PRECISION_H */
