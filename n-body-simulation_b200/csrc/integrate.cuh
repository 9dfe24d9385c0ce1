// integrate.cuh -- the kick / drift expression types of the reference's steppers (src/IntegrationMethods.cu:13-349),
// shared by the stand-alone sub-step kernel (integrate.cu) and by the force kernels that apply them in their epilogue
// (bh.cu, direct.cu: SURVEY.md section 8f-1).  A `double` literal or coefficient in the reference promotes the whole
// product to FP64 and the result is rounded once on the `+=`; each case below reproduces that expression bit for bit.
#pragma once
#include "common.cuh"

struct SubstepArgs {
    int kick, drift;
    float dt;
    double ck, cd;
};

static __device__ __forceinline__ float kick1(const SubstepArgs& s, float v, float a, float ap) {
    switch (s.kick) {
        case NB_KICK_F32:       return fmaf(s.dt, a, v);                                          // vel += dt * acc
        case NB_KICK_HALF_F64:  return (float)__fma_rn((double)s.dt * 0.5, (double)a, (double)v);  // vel += dt * 0.5 * acc
        case NB_KICK_COEF_F64:  return (float)__fma_rn(s.ck, (double)a, (double)v);                // vel += (c*dt) * acc      (PEFRL)
        case NB_KICK_DTA_F64:   return (float)__fma_rn((double)(s.dt * a), s.ck, (double)v);       // vel += dt * acc * c      (Forest-Ruth)
        case NB_KICK_MEAN_F64:  return (float)__fma_rn(0.5 * (double)s.dt, (double)(a + ap), (double)v);  // vel += 0.5*dt*(acc+prev)
        default:                return v;
    }
}
static __device__ __forceinline__ float drift1g(const SubstepArgs& s, float p, float v, float a) {
    switch (s.drift) {
        case NB_DRIFT_F32:      return fmaf(v, s.dt, p);                                           // pos += vel * dt
        case NB_DRIFT_HALF_F64: return (float)__fma_rn((double)v, (double)s.dt * 0.5, (double)p);   // pos += vel * (dt * 0.5)
        case NB_DRIFT_COEF_F64: return (float)__fma_rn(s.cd, (double)v, (double)p);                 // pos += (c*dt) * vel      (PEFRL)
        case NB_DRIFT_VDT_F64:  return (float)__fma_rn(__dmul_rn((double)(v * s.dt), s.cd), 0.5, (double)p);   // pos += vel*dt*c*0.5 (Forest-Ruth)
        case NB_DRIFT_VV_F64:   // pos += vel*dt + acc*0.5*dt*dt   (Velocity Verlet)
            return (float)((double)p + ((double)(v * s.dt) + __dmul_rn(__dmul_rn((double)a * 0.5, (double)s.dt), (double)s.dt)));
        default:                return p;
    }
}

