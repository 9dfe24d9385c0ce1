// ComputeMethods.cpp -- the reference's compute plug-ins (src/ComputeMethods.cpp:8-53) on the B200 engine.
// The only translation unit that sees both NbodySystem and the C ABI.  Compile it in place of the
// reference's ComputeMethods.cpp and link libnbody_b200.so; DirectComputeKernels.cu and
// BarnesHutKernels.cu drop out of the build.
#include <cstdio>

#include "NbodySystem.h"   // the reference's own header (unchanged)
#include "nbody_b200.h"

using namespace ComputeMethods;

ComputeMethod::ComputeMethod(int knodes, const float softening) : SOFTENING(softening), system(nullptr), knodes(knodes) {}

void ComputeMethod::createEngine(int bodies) {
    // The reference never selects a device: use whichever is current.
    int dev = 0;
    cudaGetDevice(&dev);
    if (nb_create(&engine, dev, bodies) != NB_OK) {
        // the reference prints CUDA errors and carries on (src/BarnesHutKernels.cu:15-17); a missing engine
        // cannot be carried on from, so say why and stop
        std::fprintf(stderr, "nbody_b200: %s\n", nb_last_error(nullptr));
        std::abort();
    }
}

Direct::Direct(const float softening) : ComputeMethod(1, softening) {}

void Direct::setSystem(NbodySystem* system) {
    // called from the NbodySystem constructor BEFORE the buffers exist (src/NbodySystem.cpp:15): only N is read
    this->system = system;
    createEngine(system->N);
}

void Direct::computeAcc() {
    // src/ComputeMethods.cpp:19-24: legacy default stream, asynchronous, no error check in the reference
    const int dim = system->space == R2 ? 2 : 3;
    int rc = nb_direct_accel(engine, reinterpret_cast<const float*>(system->device.pos_mass),
                             reinterpret_cast<float*>(system->device.acc), system->N, dim, SOFTENING, NB_DIRECT_AUTO,
                             nullptr);
    if (rc != NB_OK) std::printf("Direct  -  %s\n", nb_last_error(engine));
}

BarnesHut::BarnesHut(float theta, int knodes, const float softening) : ComputeMethod(knodes, softening), theta(theta) {}

void BarnesHut::setSystem(NbodySystem* system) {
    this->system = system;
    if (system->space == R3) {   // src/ComputeMethods.cpp:35-38, same message and the same silent downgrade
        printf("\n\n\nBarnes-Hut algorithm is not yet available for simulations in 3D space\nSimulating in 2D ...\n\n");
        system->space = R2;
    }
    createEngine(system->N);     // replaces the five cudaMallocs of src/ComputeMethods.cpp:40-44
}

void BarnesHut::computeAcc() {
    // src/ComputeMethods.cpp:47-53.  The reference returns fully synchronised and prints CUDA errors; the
    // engine is stream-ordered on the same (legacy default) stream, so later kernels of the integrator see the
    // accelerations without a host sync.  NB_BH_ACCUMULATE_BOUNDS would reproduce the reference's never-reset
    // bounding box; the default is the fresh box.
    int rc = nb_bh_accel(engine, reinterpret_cast<const float*>(system->device.pos_mass),
                         reinterpret_cast<float*>(system->device.acc), system->N, theta, SOFTENING, NB_BH_FRESH_BOUNDS,
                         nullptr);
    if (rc != NB_OK) printf("Barnes-Hut  -  %s\n", nb_last_error(engine));
}
