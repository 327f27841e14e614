/* quest-b200: debugging entry points of the QuEST C API.
 * Replaces reference QuEST/src/QuEST_debug.h:23-53 (same names, same argument meaning).
 * The reference's own test harness needs initStateDebug and setDensityAmps.
 */
#ifndef QUEST_DEBUG_H
#define QUEST_DEBUG_H

#include "QuEST.h"

#ifdef __cplusplus
extern "C" {
#endif

/* one qubit fixed to `outcome`, every other qubit in |+>               [ref:QuEST_debug.h:23] */
void initStateOfSingleQubit(Qureg* qureg, int qubitId, int outcome);
/* amp_k = (2k/10, (2k+1)/10) with k the GLOBAL index; unnormalised on purpose  [ref:32] */
void initStateDebug(Qureg qureg);
/* reads "re, im" lines ('#' lines skipped) into the state                      [ref:38] */
void initStateFromSingleFile(Qureg* qureg, char filename[200], QuESTEnv env);
/* 1 iff every component of every local amplitude differs by <= precision       [ref:45] */
int compareStates(Qureg mq1, Qureg mq2, qreal precision);
/* overwrite all 4^n elements of a density register                             [ref:50] */
void setDensityAmps(Qureg qureg, qreal* reals, qreal* imags);
/* declared by the reference's debug header                                     [ref:53] */
int QuESTPrecision(void);

#ifdef __cplusplus
}
#endif

#endif /* QUEST_DEBUG_H */
