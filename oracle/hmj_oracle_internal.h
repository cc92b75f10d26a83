/* internal helpers shared by the oracle translation units (TEST INFRASTRUCTURE ONLY) */
#ifndef HMJ_ORACLE_INTERNAL_H_
#define HMJ_ORACLE_INTERNAL_H_

#include "hmj_oracle.h"

#define MJ_MINVAL 1e-15
#define MJ_MAXVAL 1e10
#define MJ_PI 3.14159265358979323846
#define MJ_MINIMP 0.0001
#define MJ_MAXIMP 0.9999

double orc_normalize3(double* v);
double orc_normalize4(double* q);
void orc_mul_quat(double* res, const double* a, const double* b);
void orc_quat2mat(double* r, const double* q);
void orc_rot_vec_quat(double* res, const double* v, const double* q);
void orc_rot_vec_mat(double* res, const double* v, const double* m);
void orc_rot_vec_matT(double* res, const double* v, const double* m);
void orc_cross(double* r, const double* a, const double* b);

void orc_factor_i(const hmj_model_t* m, double* qLD, double* diaginv, int* status);
void orc_solve_ld(const hmj_model_t* m, double* x, const double* qLD, const double* diaginv);
void orc_mul_m(const hmj_model_t* m, const double* qM, double* res, const double* vec);
void orc_jac(const hmj_model_t* m, const orc_data_t* d, double* jacp, double* jacr, const double* point, int body);
void orc_reference_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);

#endif
