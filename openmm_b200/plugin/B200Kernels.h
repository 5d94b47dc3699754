// B200Kernels.h -- KernelImpl subclasses of the `B200` platform (interfaces: olla/include/openmm/kernels.h).
#pragma once
#include "B200Platform.h"
#include "openmm/kernels.h"
#include <map>
#include <string>

namespace B200Plugin {
using OpenMM::ContextImpl; using OpenMM::Platform; using OpenMM::System; using OpenMM::Vec3;
typedef B200Platform::PlatformData PData;

// kernels.h:88-126
class B200CalcForcesAndEnergyKernel : public OpenMM::CalcForcesAndEnergyKernel {
public:
    B200CalcForcesAndEnergyKernel(std::string name, const Platform& platform, PData& d) : CalcForcesAndEnergyKernel(name, platform), d(d) {}
    void initialize(const System& system) override {}
    void beginComputation(ContextImpl& context, bool includeForce, bool includeEnergy, int groups) override;
    double finishComputation(ContextImpl& context, bool includeForce, bool includeEnergy, int groups, bool& valid) override;
private:
    PData& d;
};

// kernels.h:132-246
class B200UpdateStateDataKernel : public OpenMM::UpdateStateDataKernel {
public:
    B200UpdateStateDataKernel(std::string name, const Platform& platform, PData& d) : UpdateStateDataKernel(name, platform), d(d) {}
    void initialize(const System& system) override { numParticles = system.getNumParticles(); masses.resize(numParticles); for (int i = 0; i < numParticles; i++) masses[i] = system.getParticleMass(i); }
    double getTime(const ContextImpl& context) const override;
    void setTime(ContextImpl& context, double time) override;
    long long getStepCount(const ContextImpl& context) const override;
    void setStepCount(const ContextImpl& context, long long count) override;
    void getPositions(ContextImpl& context, std::vector<Vec3>& positions, bool allowPeriodic = false) override;
    void setPositions(ContextImpl& context, const std::vector<Vec3>& positions) override;
    void getVelocities(ContextImpl& context, std::vector<Vec3>& velocities) override;
    void setVelocities(ContextImpl& context, const std::vector<Vec3>& velocities) override;
    void computeShiftedVelocities(ContextImpl& context, double timeShift, std::vector<Vec3>& velocities) override;
    void getForces(ContextImpl& context, std::vector<Vec3>& forces) override;
    void getEnergyParameterDerivatives(ContextImpl& context, std::map<std::string, double>& derivs) override {}
    void getPeriodicBoxVectors(ContextImpl& context, Vec3& a, Vec3& b, Vec3& c) const override;
    void setPeriodicBoxVectors(ContextImpl& context, const Vec3& a, const Vec3& b, const Vec3& c) override;
    void createCheckpoint(ContextImpl& context, std::ostream& stream) override;
    void loadCheckpoint(ContextImpl& context, std::istream& stream) override;
private:
    PData& d;
    int numParticles = 0;
    std::vector<double> masses;
};

// kernels.h:251-278
class B200ApplyConstraintsKernel : public OpenMM::ApplyConstraintsKernel {
public:
    B200ApplyConstraintsKernel(std::string name, const Platform& platform, PData& d) : ApplyConstraintsKernel(name, platform), d(d) {}
    void initialize(const System& system) override {}
    void apply(ContextImpl& context, double tol) override { d.all([&](b200_ctx* c, int) { return b200_apply_constraints(c, tol); }); }
    void applyToVelocities(ContextImpl& context, double tol) override { d.all([&](b200_ctx* c, int) { return b200_apply_velocity_constraints(c, tol); }); }
private:
    PData& d;
};

// kernels.h:283-302 (virtual sites are rejected at context creation)
class B200VirtualSitesKernel : public OpenMM::VirtualSitesKernel {
public:
    B200VirtualSitesKernel(std::string name, const Platform& platform) : VirtualSitesKernel(name, platform) {}
    void initialize(const System& system) override {}
    void computePositions(ContextImpl& context) override {}
};

// kernels.h:307-329: host L-BFGS of the reference driving B200 force evaluations (ReferenceMinimize.cpp:46-66)
class B200MinimizeKernel : public OpenMM::MinimizeKernel {
public:
    B200MinimizeKernel(std::string name, const Platform& platform) : MinimizeKernel(name, platform) {}
    void initialize(const System& system) override {}
    void execute(ContextImpl& context, double tolerance, int maxIterations, OpenMM::MinimizationReporter* reporter) override;
};

// kernels.h:626-688
class B200CalcNonbondedForceKernel : public OpenMM::CalcNonbondedForceKernel {
public:
    B200CalcNonbondedForceKernel(std::string name, const Platform& platform, PData& d) : CalcNonbondedForceKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::NonbondedForce& force) override;
    double execute(ContextImpl& context, bool includeForces, bool includeEnergy, bool includeDirect, bool includeReciprocal) override;
    void copyParametersToContext(ContextImpl& context, const OpenMM::NonbondedForce& force, int firstParticle, int lastParticle, int firstException, int lastException) override;
    void getPMEParameters(double& alpha, int& nx, int& ny, int& nz) const override;
    void getLJPMEParameters(double& alpha, int& nx, int& ny, int& nz) const override;
private:
    struct Params;
    void gather(const System& system, const OpenMM::NonbondedForce& force, Params& p, const std::map<std::string, double>& globals) const;
    PData& d;
    const OpenMM::NonbondedForce* forceRef = nullptr; const System* systemRef = nullptr;
    bool hasOffsets = false;
    std::map<std::string, double> globals;      // global parameters of the force and the values the device parameters were computed with
    PData::ExtraNonbonded* extra = nullptr;       // set for a second NonbondedForce of the System: it lives in a context of its own
};

// kernels.h:334-366, 408-440, 482-514
class B200CalcHarmonicBondForceKernel : public OpenMM::CalcHarmonicBondForceKernel {
public:
    B200CalcHarmonicBondForceKernel(std::string name, const Platform& platform, PData& d) : CalcHarmonicBondForceKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::HarmonicBondForce& force) override;
    double execute(ContextImpl& context, bool includeForces, bool includeEnergy) override { d.requestBonded(0, B200_GROUP_BONDS, firstTerm, numTerms); return 0.0; }
    void copyParametersToContext(ContextImpl& context, const OpenMM::HarmonicBondForce& force, int firstBond, int lastBond) override;
private:
    PData& d;
    int firstTerm = 0, numTerms = 0;
};
class B200CalcHarmonicAngleForceKernel : public OpenMM::CalcHarmonicAngleForceKernel {
public:
    B200CalcHarmonicAngleForceKernel(std::string name, const Platform& platform, PData& d) : CalcHarmonicAngleForceKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::HarmonicAngleForce& force) override;
    double execute(ContextImpl& context, bool includeForces, bool includeEnergy) override { d.requestBonded(1, B200_GROUP_ANGLES, firstTerm, numTerms); return 0.0; }
    void copyParametersToContext(ContextImpl& context, const OpenMM::HarmonicAngleForce& force, int firstAngle, int lastAngle) override;
private:
    PData& d;
    int firstTerm = 0, numTerms = 0;
};
class B200CalcPeriodicTorsionForceKernel : public OpenMM::CalcPeriodicTorsionForceKernel {
public:
    B200CalcPeriodicTorsionForceKernel(std::string name, const Platform& platform, PData& d) : CalcPeriodicTorsionForceKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::PeriodicTorsionForce& force) override;
    double execute(ContextImpl& context, bool includeForces, bool includeEnergy) override { d.requestBonded(2, B200_GROUP_TORSIONS, firstTerm, numTerms); return 0.0; }
    void copyParametersToContext(ContextImpl& context, const OpenMM::PeriodicTorsionForce& force, int firstTorsion, int lastTorsion) override;
private:
    PData& d;
    int firstTerm = 0, numTerms = 0;
};

// kernels.h:1793-1815
class B200RemoveCMMotionKernel : public OpenMM::RemoveCMMotionKernel {
public:
    B200RemoveCMMotionKernel(std::string name, const Platform& platform, PData& d) : RemoveCMMotionKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::CMMotionRemover& force) override {}
    void execute(ContextImpl& context) override { d.all([&](b200_ctx* c, int) { return b200_remove_cm_motion(c); }); }
private:
    PData& d;
};

// kernels.h:1723-1788 (MonteCarloBarostat: the Monte Carlo logic lives in MonteCarloBarostatImpl, the platform moves the molecules)
class B200ApplyMonteCarloBarostatKernel : public OpenMM::ApplyMonteCarloBarostatKernel {
public:
    B200ApplyMonteCarloBarostatKernel(std::string name, const Platform& platform, PData& d) : ApplyMonteCarloBarostatKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::Force& barostat, int components, bool rigidMolecules = true) override;
    void saveCoordinates(ContextImpl& context) override { d.all([&](b200_ctx* c, int) { return b200_barostat_save(c); }); }
    void scaleCoordinates(ContextImpl& context, double scaleX, double scaleY, double scaleZ, double scaleXY = 0, double scaleXZ = 0, double scaleYZ = 0) override {
        const double s[6] = {scaleX, scaleY, scaleZ, scaleXY, scaleXZ, scaleYZ};
        d.all([&](b200_ctx* c, int) { return b200_barostat_scale(c, s, rigid ? 1 : 0); });
    }
    void restoreCoordinates(ContextImpl& context) override { d.all([&](b200_ctx* c, int) { return b200_barostat_restore(c); }); }
    void computeKineticEnergy(ContextImpl& context, std::vector<double>& ke) override;
private:
    PData& d;
    int components = 1; bool rigid = true; int numParticles = 0;
    std::vector<double> masses;
};

// kernels.h:1392-1420
class B200IntegrateLangevinMiddleStepKernel : public OpenMM::IntegrateLangevinMiddleStepKernel {
public:
    B200IntegrateLangevinMiddleStepKernel(std::string name, const Platform& platform, PData& d) : IntegrateLangevinMiddleStepKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::LangevinMiddleIntegrator& integrator) override;
    void execute(ContextImpl& context, const OpenMM::LangevinMiddleIntegrator& integrator) override;
    double computeKineticEnergy(ContextImpl& context, const OpenMM::LangevinMiddleIntegrator& integrator) override;
private:
    PData& d;
    unsigned long long seed = 0;
};

// kernels.h:1267-1295 (leapfrog Verlet == the LangevinMiddle update with zero friction)
class B200IntegrateVerletStepKernel : public OpenMM::IntegrateVerletStepKernel {
public:
    B200IntegrateVerletStepKernel(std::string name, const Platform& platform, PData& d) : IntegrateVerletStepKernel(name, platform), d(d) {}
    void initialize(const System& system, const OpenMM::VerletIntegrator& integrator) override {}
    void execute(ContextImpl& context, const OpenMM::VerletIntegrator& integrator) override;
    double computeKineticEnergy(ContextImpl& context, const OpenMM::VerletIntegrator& integrator) override;
private:
    PData& d;
};

}  // namespace B200Plugin
