// dropin_main.cu -- a small driver written against the REFERENCE's own classes (NbodySystem, Verlet_DKD,
// DiskModel, SimulationStats).  TEST INFRASTRUCTURE.  oracle/build_dropin.sh links it twice:
//   oracle/_ref/nbody_ref_prog     with the reference's ComputeMethods.cpp + DirectComputeKernels.cu + BarnesHutKernels.cu
//   oracle/_ref/nbody_dropin_prog  with n-body-simulation_b200/adapter/ComputeMethods.cpp + libnbody_b200.so
// Everything else (NbodySystem.cpp, IntegrationMethods.cu, InitialConditions.cu, SimulationStats.cu, Callbacks.cpp)
// is the reference's source in both, so the pair shows the engine dropping in behind an unchanged caller and
// lets the energy drift of the two be compared over many steps.
//   usage: prog <direct|bh> <n> <steps> <theta> <out.bin> [verlet|pefrl|forestruth|euler_kd|euler_dk|vverlet] [disk|cold|merger|load:<file.params>] [checkpoint_dir]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <tuple>
#include <string>
#include <unistd.h>

#include "NbodySystem.h"

int main(int argc, char** argv) {
    if (argc < 6) {
        std::printf("usage: %s <direct|bh> <n> <steps> <theta> <out.bin>\n", argv[0]);
        return 2;
    }
    const bool bh = std::strcmp(argv[1], "bh") == 0;
    int n = std::atoi(argv[2]);
    const int steps = std::atoi(argv[3]);
    const float theta = (float)std::atof(argv[4]);
    const float softening = 1.01f, dt = 0.001f;   // src/main.cu:7,12

    ComputeMethods::ComputeMethod* compute =
        bh ? (ComputeMethods::ComputeMethod*)new ComputeMethods::BarnesHut(theta, 3, softening)
           : (ComputeMethods::ComputeMethod*)new ComputeMethods::Direct(softening);
    const char* integ = argc > 6 ? argv[6] : "verlet";
    const char* icname = argc > 7 ? argv[7] : "disk";
    IntegrationMethods::IntegrationMethod* integrator;
    if (!std::strcmp(integ, "pefrl")) integrator = new IntegrationMethods::PEFRL();                 // 4 force evaluations per step
    else if (!std::strcmp(integ, "forestruth")) integrator = new IntegrationMethods::ForestRuth();  // 3
    else if (!std::strcmp(integ, "euler_kd")) integrator = new IntegrationMethods::Euler_Symplectic_KD();
    else if (!std::strcmp(integ, "euler_dk")) integrator = new IntegrationMethods::Euler_Symplectic_DK();
    else if (!std::strcmp(integ, "vverlet")) integrator = new IntegrationMethods::Velocity_Verlet();
    else integrator = new IntegrationMethods::Verlet_DKD();                                          // "Leapfrog"
    const bool load = !std::strncmp(icname, "load:", 5);
    // either the plain constructor + an initial-condition generator, or the reference's own checkpoint loader
    // (src/NbodySystem.cpp:27-70), which takes N from the file size
    NbodySystem* sysp = load ? new NbodySystem(std::string(icname + 5), Space::R2, compute, integrator)
                             : new NbodySystem(n, Space::R2, compute, integrator);
    NbodySystem& system = *sysp;
    if (argc > 8) {   // run inside <checkpoint_dir> and let the reference's CheckpointSaver write configs/run_<steps>.params
        if (chdir(argv[8]) != 0) return 3;
        system.addCallback(new Callbacks::CheckpointSaver(steps, "run"));
    }
    if (load) {
    } else if (!std::strcmp(icname, "cold")) {          // cold collapse: the disc at rest (BASELINE.json config 4)
        auto ic = new InitialConditions::Modules::UniformEllipsoid({0, 0}, {0, 0}, {30000, 30000}, {70000, 70000}, {0, 0});
        system.initSystem(0, n, ic);
    } else if (!std::strcmp(icname, "merger")) { // two discs on a collision course (config 5)
        // (UniformEllipsoid, not DiskModel: DiskModel evaluates forces while the other half is still uninitialised,
        //  i.e. n/2 coincident massless bodies at the origin, which the reference's tree build does not survive)
        auto a = new InitialConditions::Modules::UniformEllipsoid({-60000, -15000}, {40, 0}, {30000, 30000}, {70000, 70000}, {0, 0});
        auto b = new InitialConditions::Modules::UniformEllipsoid({60000, 15000}, {-40, 0}, {30000, 30000}, {70000, 70000}, {0, 0});
        system.initSystem(0, n / 2, a);
        system.initSystem(n / 2, n - n / 2, b);
    } else {
        auto ic = new InitialConditions::DiskModel({0, 0}, {0, 0}, {30000, 30000}, {70000, 70000}, {0, 0});   // src/main.cu:26-32
        system.initSystem(0, n, ic);
    }

    n = system.N;
    auto [m0, k0, p0] = SimulationStats::computeStats(system.space, system.device.pos_mass, system.device.vel, n);
    system.simulate(steps, dt);
    cudaDeviceSynchronize();
    auto [m1, k1, p1] = SimulationStats::computeStats(system.space, system.device.pos_mass, system.device.vel, n);
    std::printf("RESULT method=%s n=%d steps=%d E0=%.12e E1=%.12e drift=%.6e moment_drift=%.6e\n", argv[1], n, steps, k0 + p0,
                k1 + p1, (k1 + p1 - k0 - p0) / (k0 + p0), (m1 - m0) / m0);
    // the three sums of SimulationStats::computeStats (src/SimulationStats.cu:7-27) before and after, one by one
    std::printf("STATS M0=%.15e K0=%.15e P0=%.15e M1=%.15e K1=%.15e P1=%.15e\n", m0, k0, p0, m1, k1, p1);
    system.updateHostData();
    FILE* f = std::fopen(argv[5], "wb");
    if (f) {
        std::fwrite(system.host.pos_mass, sizeof(float4), n, f);
        std::fwrite(system.host.vel, sizeof(float4), n, f);
        std::fclose(f);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) std::printf("CUDA error: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}
