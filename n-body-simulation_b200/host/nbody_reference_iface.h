// nbody_reference_iface.h -- the types CudaNaiveSimulation is written against.
//
// When the reference's headers are on the include path (-I<reference>/include) they are used as
// they are, so the class derives from the reference's own SimulationAlgorithm and can be handed to
// its NBodySimulation unchanged (include/Simulation.h:14).  When they are not (stand-alone builds
// and the GPU test box, where the reference tree does not exist), the declarations below stand in
// for them.  They restate only the INTERFACE of include/Particle.hpp:3-59 and
// include/SimulationAlgorithm.h:4-22 -- field order, types, defaults, the virtual signature --
// which is what layout- and call-compatibility needs; none of the reference's operator code.
#pragma once
#include <vector>

#if defined(__has_include) && __has_include("SimulationAlgorithm.h") && !defined(NBODY_STANDALONE_IFACE)
#include "SimulationAlgorithm.h"
#define NBODY_USING_REFERENCE_HEADERS 1
#else
#define NBODY_USING_REFERENCE_HEADERS 0

struct vector3_t {          // three packed floats (Particle.hpp:3-7)
    float x, y, z;
};

struct particle_t {         // 40 bytes, offsets 0/12/24/36 (Particle.hpp:53-59)
    vector3_t position;
    vector3_t velocity;
    vector3_t acceleration;
    float mass;
};

struct SimParams {          // SimulationAlgorithm.h:4-9
    float G = 1.0f;
    float dt = 0.05f;
    float min_r2 = 1e-8f;   // softening^2, added to r^2
};

class SimulationAlgorithm { // SimulationAlgorithm.h:18-22
public:
    virtual ~SimulationAlgorithm() = default;
    virtual void computeStep(std::vector<particle_t> &particles, const SimParams &params) = 0;
};
#endif

static_assert(sizeof(particle_t) == 40, "particle_t must be the reference's 40-byte record");
static_assert(sizeof(SimParams) == 12, "SimParams must be three floats");
