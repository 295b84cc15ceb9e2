// lmpc_aux.cu — placeholder translation unit for the auxiliary batched kernels (kinematics /
// torque map, FSM update); filled in by later sections of the build.
#include "lmpc_internal.h"
