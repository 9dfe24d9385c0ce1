// ComputeMethods.h -- drop-in replacement for the reference's src/ComputeMethods.h.
//
// Same namespace, class names, constructor signatures, public `knodes` and virtual surface as the
// reference (src/ComputeMethods.h:8-42), so NbodySystem, IntegrationMethods and InitialConditions compile
// against it unchanged.  The private members differ: the reference's BarnesHut owns five raw scratch
// pointers (src/ComputeMethods.h:33-35); here both methods own one engine handle of the C ABI
// (include/nbody_b200.h) and the scratch lives behind it.
#pragma once
#include <cuda_runtime.h>

class NbodySystem;
struct nb_context;

namespace ComputeMethods {

class ComputeMethod {
protected:
    const float SOFTENING;
    NbodySystem* system;
    nb_context* engine = nullptr;
    void createEngine(int bodies);

public:
    int knodes = 1;

    ComputeMethod(int knodes, const float softening);

    virtual void setSystem(NbodySystem* system) = 0;
    virtual void computeAcc() = 0;
};

class Direct : public ComputeMethod {
public:
    Direct(const float softening);

    void setSystem(NbodySystem* system);
    void computeAcc();
};

class BarnesHut : public ComputeMethod {
    float theta;

public:
    BarnesHut(float theta, int knodes, const float softening);

    void setSystem(NbodySystem* system);
    void computeAcc();
};

}  // namespace ComputeMethods
