/*
 * oracle/teach_oracle.c -- CPU restatement of the reference's TEACH-T sudden-expansion ("backward step") solver,
 * laminar subset: calcu / calcv / calcp with their boundary hooks modu / modv and the line solver.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/cavity_oracle.c header for the rule).
 *
 * Sources restated (paths under /root/reference/backward-step-flow-2d/):
 *   orc_teach_init      <- init      general.f90:69-140
 *   orc_teach_grid      <- contro    control.f90:72-90 (x stretching factor epsx made a parameter)
 *   orc_teach_initial   <- init / contro  general.f90:142-161, control.f90:148-179,197-203
 *   orc_teach_lisolv    <- lisolv    general.f90:294-327
 *   modu / modv         <- promod.f90:26-95, 99-133 (laminar branch of the wall functions: yplus = 11 <= 11.63)
 *   orc_teach_assemble_u/v/p, correct <- calc.f90:20-124, 137-220, 231-304
 *
 * The reference computes in default REAL (FP32); BASELINE.json:north_star asks for FP64.  The body
 * (teach_body.inc) is therefore compiled twice: *_f64 is the parity oracle of the FP64 CUDA path, *_f32 is the
 * reference "as written", used to check that the two agree to single-precision accuracy.
 * k-epsilon, energy equation and props: teach_turb.inc, same two precisions.
 *
 * PARITY PIN: none exists -- the reference ships no outputs for this program and SURVEY 8(c) has no table for
 * it; "parity unpinned" applies to this family.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define X1(a, i) ((a)[(i) - 1])

#define REAL double
#define NAME(x) x##_f64
#define SQRT sqrt
#define POW pow
#define LOG log
#define EXP exp
#define RABS(x) fabs(x)
#define RMAX(a, b) ((a) > (b) ? (a) : (b))
#include "teach_body.inc"
#include "teach_turb.inc"
#undef REAL
#undef NAME
#undef SQRT
#undef POW
#undef LOG
#undef EXP
#undef RABS

#define REAL float
#define NAME(x) x##_f32
#define SQRT sqrtf
#define POW powf
#define LOG logf
#define EXP expf
#define RABS(x) fabsf(x)
#include "teach_body.inc"
#include "teach_turb.inc"
