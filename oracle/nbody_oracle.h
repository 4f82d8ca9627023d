/* oracle/nbody_oracle.h — TEST INFRASTRUCTURE ONLY (see nbody_oracle.c). */
#ifndef NBODY_ORACLE_H
#define NBODY_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif
#define NBO_FORCE_PAIRWISE 0
#define NBO_FORCE_PAIRWISE_AVX2_FP32 1
#define NBO_FORCE_ADAPTIVE_MUTUAL 2
#define NBO_INTEG_EULER 0
#define NBO_INTEG_VELOCITY_VERLET 1

void   nbo_set_threads(int t);
int    nbo_max_threads(void);
double nbo_average_mass(long n, const double* posMass4);
int    nbo_forces(long n, const double* posMass4, int force_method, double* accel3);
int    nbo_forces_subset(long n, const double* posMass4, int force_method,
                         long n_idx, const long* idx, double* accel3);
void   nbo_com_velocity(long n, const double* posMass4, const double* vel4, double vcm[3]);
int    nbo_step(long n, double* posMass4, double* vel4, int integ, int force_method, double dt, int steps);
double nbo_energy(long n, const double* posMass4, const double* vel4);
void   nbo_momentum(long n, const double* posMass4, const double* vel4, double p[3]);
#ifdef __cplusplus
}
#endif
#endif
