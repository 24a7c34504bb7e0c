// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C harness around the UNMODIFIED reference CPU classes of HOOMD-blue 2.3.0, compiled from the
// sources where they lie under /root/reference (see oracle/Makefile, target _ref/libhoomd_ref.so).
// It drives the reference's own CellList / NeighborListBinned / PotentialPairLJ / IntegratorTwoStep /
// TwoStepNVE / TwoStepLangevin / SFCPackUpdater / ComputeThermo objects exactly as the reference's unit
// tests do (hoomd/md/test/test_lj_force.cc:51-70, hoomd/md/test/test_nve_integrator.cc:262-300) and
// exposes their arrays through a flat C API so that tests/ and bench.py (cpu_baseline / --impl reference)
// can compare the CUDA path and the C restatement (oracle/oracle.c) against the real thing.
//
// Nothing in hoomd_b200/ may link or load this file.

#include "hoomd/ExecutionConfiguration.h"
#include "hoomd/SystemDefinition.h"
#include "hoomd/ParticleData.h"
#include "hoomd/ParticleGroup.h"
#include "hoomd/CellList.h"
#include "hoomd/Variant.h"
#include "hoomd/SFCPackUpdater.h"
#include "hoomd/ComputeThermo.h"
#include "hoomd/md/TwoStepNVTMTK.h"
#include "hoomd/md/TwoStepNPTMTK.h"
#include "hoomd/BoxResizeUpdater.h"
#include "hoomd/Saru.h"
#include "hoomd/md/NeighborListBinned.h"
#include "hoomd/md/AllPairPotentials.h"
#include "hoomd/md/AllBondPotentials.h"
#include "hoomd/md/IntegratorTwoStep.h"
#include "hoomd/md/TwoStepNVE.h"
#include "hoomd/md/TwoStepLangevin.h"

#include <memory>
#include <vector>
#include <cstring>
#include <chrono>

namespace
{
struct RefSystem
    {
    std::shared_ptr<ExecutionConfiguration> exec_conf;
    std::shared_ptr<SystemDefinition> sysdef;
    std::shared_ptr<ParticleData> pdata;
    std::shared_ptr<ParticleGroup> group_all;
    std::shared_ptr<ParticleGroup> group_method; // the group the integration method acts on (default: all)
    std::shared_ptr<CellList> cl;
    std::shared_ptr<NeighborListBinned> nlist;
    std::shared_ptr<ForceCompute> lj;     // the pair force: PotentialPairLJ or another PotentialPair<evaluator>
    std::shared_ptr<ForceCompute> bond;   // PotentialBondFENE or PotentialBondHarmonic (md.bond.fene / md.bond.harmonic)
    std::shared_ptr<IntegratorTwoStep> integrator;
    std::shared_ptr<TwoStepNVE> nve;
    std::shared_ptr<TwoStepLangevin> langevin;
    std::shared_ptr<SFCPackUpdater> sorter;
    unsigned int check_period = 1;
    std::shared_ptr<ComputeThermo> thermo;
    std::shared_ptr<ComputeThermo> thermo_nvt; // the method's own compute (hoomd/md/integrate.py:370-372)
    std::shared_ptr<TwoStepNVTMTK> nvt;
    std::shared_ptr<ComputeThermo> thermo_npt, thermo_npt_t; // half-step / full-step computes of md.integrate.npt
    std::shared_ptr<TwoStepNPTMTK> npt;
    unsigned int sorter_period = 0;
    unsigned int sorter_next = 0;
    unsigned int timestep = 0;
    bool virial = false;
    std::string error;
    };

void apply_flags(RefSystem* s)
    {
    PDataFlags flags(0);
    flags[pdata_flag::potential_energy] = 1;
    if (s->virial)
        {
        flags[pdata_flag::pressure_tensor] = 1;
        flags[pdata_flag::isotropic_virial] = 1;
        }
    s->pdata->setFlags(flags);
    }
} // namespace

namespace
{
inline void make_param(Scalar2& p, const float* q) { p = make_scalar2(q[0], q[1]); }
inline void make_param(Scalar4& p, const float* q) { p = make_scalar4(q[0], q[1], q[2], q[3]); }

template<class P>
std::shared_ptr<ForceCompute> make_pair(RefSystem* s, const float* rcut, const float* params4, const float* ron,
                                        int shift_mode)
    {
    unsigned int nt = s->pdata->getNTypes();
    std::shared_ptr<P> f(new P(s->sysdef, s->nlist));
    for (unsigned int i = 0; i < nt; i++)
        for (unsigned int j = i; j < nt; j++)
            {
            typename P::param_type prm;
            make_param(prm, params4 + 4 * (i * nt + j));
            f->setParams(i, j, prm);
            f->setRcut(i, j, rcut[i * nt + j]);
            if (ron) f->setRon(i, j, ron[i * nt + j]);
            s->nlist->setRCutPair(i, j, rcut[i * nt + j]);
            }
    f->setShiftMode((typename P::energyShiftMode)shift_mode);
    return f;
    }
} // namespace

#define REF_TRY(s) try {
#define REF_CATCH(s, rv) } catch (const std::exception& e) { (s)->error = e.what(); return rv; }

extern "C"
{

//! Create a system of N particles of \a ntypes types in an orthorhombic periodic box (tilts 0)
void* ref_create(unsigned int N, float Lx, float Ly, float Lz, unsigned int ntypes)
    {
    RefSystem* s = new RefSystem;
    try
        {
        std::shared_ptr<Messenger> msg(new Messenger());
        msg->setNoticeLevel(0);
        s->exec_conf = std::shared_ptr<ExecutionConfiguration>(
            new ExecutionConfiguration(ExecutionConfiguration::CPU, -1, false, false, msg));
        BoxDim box(Lx, Ly, Lz);
        s->sysdef = std::shared_ptr<SystemDefinition>(new SystemDefinition(N, box, ntypes, 0, 0, 0, 0, s->exec_conf));
        s->pdata = s->sysdef->getParticleData();
        std::shared_ptr<ParticleSelector> sel(new ParticleSelectorTag(s->sysdef, 0, N - 1));
        s->group_all = std::shared_ptr<ParticleGroup>(new ParticleGroup(s->sysdef, sel));
        apply_flags(s);
        }
    catch (const std::exception& e)
        {
        s->error = e.what();
        }
    return s;
    }

//! Same with \a n_bond_types bond types (BondData is sized at SystemDefinition construction)
void* ref_create_bonded(unsigned int N, float Lx, float Ly, float Lz, unsigned int ntypes, unsigned int n_bond_types)
    {
    RefSystem* s = new RefSystem;
    try
        {
        std::shared_ptr<Messenger> msg(new Messenger());
        msg->setNoticeLevel(0);
        s->exec_conf = std::shared_ptr<ExecutionConfiguration>(
            new ExecutionConfiguration(ExecutionConfiguration::CPU, -1, false, false, msg));
        BoxDim box(Lx, Ly, Lz);
        s->sysdef = std::shared_ptr<SystemDefinition>(
            new SystemDefinition(N, box, ntypes, n_bond_types, 0, 0, 0, s->exec_conf));
        s->pdata = s->sysdef->getParticleData();
        std::shared_ptr<ParticleSelector> sel(new ParticleSelectorTag(s->sysdef, 0, N - 1));
        s->group_all = std::shared_ptr<ParticleGroup>(new ParticleGroup(s->sysdef, sel));
        apply_flags(s);
        }
    catch (const std::exception& e)
        {
        s->error = e.what();
        }
    return s;
    }

//! BondData::addBondedGroup for every bond (tags2 = pairs of particle tags)
int ref_add_bonds(void* h, unsigned int nbonds, const unsigned int* tags2, const unsigned int* types)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<BondData> bd = s->sysdef->getBondData();
    for (unsigned int b = 0; b < nbonds; b++)
        bd->addBondedGroup(Bond(types[b], tags2[2 * b], tags2[2 * b + 1]));
    REF_CATCH(s, 1)
    return 0;
    }

//! md.bond.fene (kind 0: params (K, r0, lj1, lj2)) or md.bond.harmonic (kind 1: params (K, r0)) per bond type
int ref_setup_bond(void* h, int kind, const float* K, const float* r0, const float* lj1, const float* lj2)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    unsigned int nb = s->sysdef->getBondData()->getNTypes();
    if (kind == 0)
        {
        std::shared_ptr<PotentialBondFENE> f(new PotentialBondFENE(s->sysdef, ""));
        for (unsigned int t = 0; t < nb; t++) f->setParams(t, make_scalar4(K[t], r0[t], lj1[t], lj2[t]));
        s->bond = f;
        }
    else
        {
        std::shared_ptr<PotentialBondHarmonic> f(new PotentialBondHarmonic(s->sysdef, ""));
        for (unsigned int t = 0; t < nb; t++) f->setParams(t, make_scalar2(K[t], r0[t]));
        s->bond = f;
        }
    if (s->integrator) s->integrator->addForceCompute(s->bond);
    REF_CATCH(s, 1)
    return 0;
    }

//! nlist.reset_exclusions(['bond']) (on) / reset_exclusions([]) (off)
int ref_nlist_exclusions_from_bonds(void* h, int on)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->nlist->clearExclusions();
    if (on) s->nlist->addExclusionsFromBonds();
    s->nlist->forceUpdate();
    REF_CATCH(s, 1)
    return 0;
    }

void ref_destroy(void* h)
    {
    delete static_cast<RefSystem*>(h);
    }

const char* ref_last_error(void* h)
    {
    return static_cast<RefSystem*>(h)->error.c_str();
    }

//! Non-periodic directions (used to emulate a domain-decomposed local box)
void ref_set_periodic(void* h, int px, int py, int pz)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    BoxDim box = s->pdata->getGlobalBox();
    box.setPeriodic(make_uchar3(px, py, pz));
    s->pdata->setGlobalBox(box);
    }

//! Set particle state in index order; tags are the identity at this point (tag == index)
void ref_set_particles(void* h, const float* pos4, const float* vel4, const int* image3)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int N = s->pdata->getN();
    ArrayHandle<Scalar4> h_pos(s->pdata->getPositions(), access_location::host, access_mode::readwrite);
    ArrayHandle<Scalar4> h_vel(s->pdata->getVelocities(), access_location::host, access_mode::readwrite);
    ArrayHandle<int3> h_img(s->pdata->getImages(), access_location::host, access_mode::readwrite);
    if (pos4) memcpy(h_pos.data, pos4, sizeof(Scalar4) * N);
    if (vel4) memcpy(h_vel.data, vel4, sizeof(Scalar4) * N);
    if (image3) memcpy(h_img.data, image3, sizeof(int3) * N);
    if (s->nlist) s->nlist->forceUpdate();
    }

void ref_set_accel(void* h, const float* accel3)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    ArrayHandle<Scalar3> h_accel(s->pdata->getAccelerations(), access_location::host, access_mode::readwrite);
    memcpy(h_accel.data, accel3, sizeof(Scalar3) * s->pdata->getN());
    s->pdata->notifyAccelSet();
    }

//! md.nlist.cell + md.pair.lj
/*! \param rcut, lj1, lj2 ntypes x ntypes matrices (row-major, symmetric)
    \param shift_mode 0 none, 1 shift, 2 xplor (hoomd/md/PotentialPair.h:109-114)
    \param full_storage 1 = full neighbor list (what every GPU class forces), 0 = half (CPU default) */
int ref_setup_lj(void* h, float r_buff, int check_period, int full_storage, const float* rcut,
                 const float* lj1, const float* lj2, const float* ron, int shift_mode)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    unsigned int nt = s->pdata->getNTypes();
    s->cl = std::shared_ptr<CellList>(new CellList(s->sysdef));
    s->nlist = std::shared_ptr<NeighborListBinned>(new NeighborListBinned(s->sysdef, Scalar(0.0), r_buff, s->cl));
    s->check_period = check_period;
    s->nlist->setEvery(check_period, true);
    s->nlist->setStorageMode(full_storage ? NeighborList::full : NeighborList::half);
    std::shared_ptr<PotentialPairLJ> lj(new PotentialPairLJ(s->sysdef, s->nlist));
    for (unsigned int i = 0; i < nt; i++)
        for (unsigned int j = i; j < nt; j++)
            {
            lj->setParams(i, j, make_scalar2(lj1[i * nt + j], lj2[i * nt + j]));
            lj->setRcut(i, j, rcut[i * nt + j]);
            if (ron) lj->setRon(i, j, ron[i * nt + j]);
            s->nlist->setRCutPair(i, j, rcut[i * nt + j]);
            }
    lj->setShiftMode((PotentialPairLJ::energyShiftMode)shift_mode);
    s->lj = lj;
    REF_CATCH(s, 1)
    return 0;
    }


//! md.nlist.cell + md.pair.<evaluator>: 0 lj, 1 gauss, 2 yukawa, 3 mie, 4 force_shifted_lj, 5 morse
/*! \param params4 ntypes x ntypes x 4 floats: the evaluator's param_type padded to four components */
int ref_setup_pair(void* h, int evaluator, float r_buff, int check_period, int full_storage, const float* rcut,
                   const float* params4, const float* ron, int shift_mode)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->cl = std::shared_ptr<CellList>(new CellList(s->sysdef));
    s->nlist = std::shared_ptr<NeighborListBinned>(new NeighborListBinned(s->sysdef, Scalar(0.0), r_buff, s->cl));
    s->check_period = check_period;
    s->nlist->setEvery(check_period, true);
    s->nlist->setStorageMode(full_storage ? NeighborList::full : NeighborList::half);
    switch (evaluator)
        {
        case 0: s->lj = make_pair<PotentialPairLJ>(s, rcut, params4, ron, shift_mode); break;
        case 1: s->lj = make_pair<PotentialPairGauss>(s, rcut, params4, ron, shift_mode); break;
        case 2: s->lj = make_pair<PotentialPairYukawa>(s, rcut, params4, ron, shift_mode); break;
        case 3: s->lj = make_pair<PotentialPairMie>(s, rcut, params4, ron, shift_mode); break;
        case 4: s->lj = make_pair<PotentialPairForceShiftedLJ>(s, rcut, params4, ron, shift_mode); break;
        case 5: s->lj = make_pair<PotentialPairMorse>(s, rcut, params4, ron, shift_mode); break;
        case 6: s->lj = make_pair<PotentialPairDPD>(s, rcut, params4, ron, shift_mode); break;
        case 7: s->lj = make_pair<PotentialPairDPDThermoDPD>(s, rcut, params4, ron, shift_mode); break; // md.pair.dpd
        default: throw std::runtime_error("unknown pair evaluator");
        }
    REF_CATCH(s, 1)
    return 0;
    }

//! md.pair.dpd(kT, seed): temperature and seed of the DPD thermostat created by ref_setup_pair(evaluator = 7)
//! (PotentialPairDPDThermo::setT / setSeed, hoomd/md/PotentialPairDPDThermo.h:60-70)
int ref_set_dpd_thermostat(void* h, float kT, unsigned int seed)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<PotentialPairDPDThermoDPD> f = std::dynamic_pointer_cast<PotentialPairDPDThermoDPD>(s->lj);
    if (!f) throw std::runtime_error("ref_set_dpd_thermostat: the pair force is not PotentialPairDPDThermoDPD");
    f->setSeed(seed);
    f->setT(std::shared_ptr<Variant>(new VariantConst(kT)));
    REF_CATCH(s, 1)
    return 0;
    }

void ref_set_virial(void* h, int on)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    s->virial = on;
    apply_flags(s);
    }

int ref_setup_integrator(void* h, float dt)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->integrator = std::shared_ptr<IntegratorTwoStep>(new IntegratorTwoStep(s->sysdef, dt));
    if (s->lj) s->integrator->addForceCompute(s->lj);
    if (s->bond) s->integrator->addForceCompute(s->bond);
    s->thermo = std::shared_ptr<ComputeThermo>(new ComputeThermo(s->sysdef, s->group_all, ""));
    REF_CATCH(s, 1)
    return 0;
    }

int ref_add_nve(void* h, int limit, float limit_val, int zero_force)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->nve = std::shared_ptr<TwoStepNVE>(new TwoStepNVE(s->sysdef, s->group_method ? s->group_method : s->group_all, true));
    if (limit) s->nve->setLimit(limit_val);
    s->nve->setZeroForce(zero_force);
    s->integrator->removeAllIntegrationMethods();
    s->integrator->addIntegrationMethod(s->nve);
    REF_CATCH(s, 1)
    return 0;
    }

int ref_add_langevin(void* h, float kT, unsigned int seed, const float* gamma, int noiseless_t)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<Variant> T(new VariantConst(kT));
    s->langevin = std::shared_ptr<TwoStepLangevin>(
        new TwoStepLangevin(s->sysdef, s->group_method ? s->group_method : s->group_all, T, seed, false, Scalar(0.0),
                            noiseless_t, false, ""));
    if (gamma)
        for (unsigned int t = 0; t < s->pdata->getNTypes(); t++)
            s->langevin->setGamma(t, gamma[t]);
    s->integrator->removeAllIntegrationMethods();
    s->integrator->addIntegrationMethod(s->langevin);
    REF_CATCH(s, 1)
    return 0;
    }

//! md.integrate.langevin(group, kT, seed, dscale=lambda, tally) (hoomd/md/integrate.py:844-940): gamma = lambda * diameter
//! when use_lambda, the reservoir energy is tallied when tally (TwoStepLangevin::setTally)
int ref_add_langevin_ex(void* h, float kT, unsigned int seed, const float* gamma, int noiseless_t, int use_lambda,
                        float lambda, int tally)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<Variant> T(new VariantConst(kT));
    s->langevin = std::shared_ptr<TwoStepLangevin>(
        new TwoStepLangevin(s->sysdef, s->group_method ? s->group_method : s->group_all, T, seed, use_lambda != 0,
                            Scalar(lambda), noiseless_t, false, ""));
    if (gamma)
        for (unsigned int t = 0; t < s->pdata->getNTypes(); t++)
            s->langevin->setGamma(t, gamma[t]);
    s->langevin->setTally(tally != 0);
    s->integrator->removeAllIntegrationMethods();
    s->integrator->addIntegrationMethod(s->langevin);
    REF_CATCH(s, 1)
    return 0;
    }

//! the "langevin_reservoir_energy" log quantity (TwoStepLangevin::getLogValue, hoomd/md/TwoStepLangevin.cc:61-73)
double ref_get_langevin_reservoir_energy(void* h)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    bool my = false;
    return s->langevin ? (double)s->langevin->getLogValue("langevin_reservoir_energy", s->timestep, my) : 0.0;
    }

//! per-particle diameters (ParticleData::getDiameters), index order
void ref_set_diameters(void* h, const float* diameter)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    ArrayHandle<Scalar> d(s->pdata->getDiameters(), access_location::host, access_mode::overwrite);
    for (unsigned int i = 0; i < s->pdata->getN(); i++) d.data[i] = diameter[i];
    }

//! hoomd.group.tag_list: the integration methods created AFTER this call act on these member tags
//! (ParticleGroup(sysdef, member_tags), hoomd/ParticleGroup.cc); n = 0 restores group.all()
int ref_set_group_tags(void* h, const unsigned int* tags, unsigned int n)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    if (n == 0)
        s->group_method.reset();
    else
        {
        std::vector<unsigned int> members(tags, tags + n);
        s->group_method = std::shared_ptr<ParticleGroup>(new ParticleGroup(s->sysdef, members));
        }
    REF_CATCH(s, 1)
    return 0;
    }

//! md.integrate.nvt(group=all, kT, tau) (hoomd/md/integrate.py:349-383): the method gets its own ComputeThermo whose NDOF
//! the integrator sets before the run (hoomd/md/integrate.py:152-160)
int ref_add_nvt(void* h, float kT, float tau)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<Variant> T(new VariantConst(kT));
    std::shared_ptr<ParticleGroup> grp = s->group_method ? s->group_method : s->group_all;
    s->thermo_nvt = std::shared_ptr<ComputeThermo>(new ComputeThermo(s->sysdef, grp, "nvt"));
    s->nvt = std::shared_ptr<TwoStepNVTMTK>(new TwoStepNVTMTK(s->sysdef, grp, s->thermo_nvt, tau, T, ""));
    s->integrator->removeAllIntegrationMethods();
    s->integrator->addIntegrationMethod(s->nvt);
    s->thermo_nvt->setNDOF(s->integrator->getNDOF(grp));
    REF_CATCH(s, 1)
    return 0;
    }

//! md.integrate.npt(group, kT, tau, P, tauP, couple, x, y, z, nph, rescale_all, gamma) (hoomd/md/integrate.py:406-506):
//! TwoStepNPTMTK with a half-step ComputeThermo of its own and a full-step one, both given the integrator's NDOF
//! (hoomd/md/integrate.py:152-160); the method requests the pressure tensor (TwoStepNPTMTK.h:152-163).
//! couple: 0 none, 1 xy, 2 xz, 3 yz, 4 xyz; flags: baro_x 1 | baro_y 2 | baro_z 4 (the box stays orthorhombic).
int ref_add_npt(void* h, float kT, float tau, float P, float tauP, int couple, unsigned int flags, int nph, int rescale_all,
                float gamma)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    if (flags & ~7u) throw std::runtime_error("ref_add_npt: tilt degrees of freedom are outside the harness");
    std::shared_ptr<Variant> T(new VariantConst(kT));
    std::shared_ptr<Variant> Pv(new VariantConst(P));
    std::shared_ptr<ParticleGroup> grp = s->group_method ? s->group_method : s->group_all;
    s->thermo_npt = std::shared_ptr<ComputeThermo>(new ComputeThermo(s->sysdef, grp, "npt"));
    s->thermo_npt_t = std::shared_ptr<ComputeThermo>(new ComputeThermo(s->sysdef, grp, "npt_t"));
    s->npt = std::shared_ptr<TwoStepNPTMTK>(new TwoStepNPTMTK(s->sysdef, grp, s->thermo_npt, s->thermo_npt_t, tau, tauP, T, Pv,
                                                              (TwoStepNPTMTK::couplingMode)couple, flags, nph != 0));
    s->npt->setRescaleAll(rescale_all != 0);
    s->npt->setGamma(gamma);
    s->integrator->removeAllIntegrationMethods();
    s->integrator->addIntegrationMethod(s->npt);
    s->thermo_npt->setNDOF(s->integrator->getNDOF(grp));
    s->thermo_npt_t->setNDOF(s->integrator->getNDOF(grp));
    s->virial = true;
    apply_flags(s);
    REF_CATCH(s, 1)
    return 0;
    }

//! the method's IntegratorVariables {eta, xi, nuxx, nuxy, nuxz, nuyy, nuyz, nuzz, xi_rot, eta_rot} and its two log
//! quantities {npt_thermostat_energy, npt_barostat_energy} (hoomd/md/TwoStepNPTMTK.cc:593-636)
int ref_get_npt_state(void* h, float* out12)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<IntegratorData> idata = s->sysdef->getIntegratorData();
    for (int k = 0; k < 12; k++) out12[k] = 0.0f;
    for (unsigned int i = 0; i < idata->getNumIntegrators(); i++)
        {
        const IntegratorVariables& v = idata->getIntegratorVariables(i);
        if (v.type == "npt_mtk" && v.variable.size() >= 10)
            for (int k = 0; k < 10; k++) out12[k] = v.variable[k];
        }
    bool mine = false;
    out12[10] = s->npt->getLogValue("npt_thermostat_energy", s->timestep, mine);
    out12[11] = s->npt->getLogValue("npt_barostat_energy", s->timestep, mine);
    REF_CATCH(s, 1)
    return 0;
    }

//! the full-step compute's pressure tensor {xx, xy, xz, yy, yz, zz}, pressure and translational kinetic energy at ts
int ref_get_pressure_tensor(void* h, unsigned int ts, float* out8)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<ComputeThermo> t = s->thermo_npt_t ? s->thermo_npt_t : s->thermo;
    t->compute(ts);
    PressureTensor P = t->getPressureTensor();
    out8[0] = P.xx; out8[1] = P.xy; out8[2] = P.xz; out8[3] = P.yy; out8[4] = P.yz; out8[5] = P.zz;
    out8[6] = t->getPressure();
    out8[7] = t->getTranslationalKineticEnergy();
    REF_CATCH(s, 1)
    return 0;
    }

//! thermostat state {xi, eta} (IntegratorVariables of the method, hoomd/md/TwoStepNVTMTK.cc:387-404)
int ref_get_nvt_state(void* h, float* xi_eta)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    // the method's accessor is protected: find its record in the IntegratorData (the last "nvt_mtk" registered)
    std::shared_ptr<IntegratorData> idata = s->sysdef->getIntegratorData();
    xi_eta[0] = xi_eta[1] = 0.0f;
    for (unsigned int i = 0; i < idata->getNumIntegrators(); i++)
        {
        const IntegratorVariables& v = idata->getIntegratorVariables(i);
        if (v.type == "nvt_mtk" && v.variable.size() >= 2)
            {
            xi_eta[0] = v.variable[0];
            xi_eta[1] = v.variable[1];
            }
        }
    REF_CATCH(s, 1)
    return 0;
    }

//! One BoxResizeUpdater::update (hoomd/BoxResizeUpdater.cc:60-136) towards the box (Lx, Ly, Lz), tilts unchanged: what
//! hoomd.update.box_resize does at a period boundary (the variants evaluated by the caller)
int ref_set_box(void* h, float Lx, float Ly, float Lz, int scale_particles)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    // updaters run in the order they were added (System::run): the sorter comes first
    if (s->sorter && s->sorter_period && s->timestep == s->sorter_next)
        {
        s->sorter->update(s->timestep);
        s->sorter_next += s->sorter_period;
        }
    BoxDim cur = s->pdata->getGlobalBox();
    std::shared_ptr<Variant> vx(new VariantConst(Lx)), vy(new VariantConst(Ly)), vz(new VariantConst(Lz));
    std::shared_ptr<Variant> xy(new VariantConst(cur.getTiltFactorXY())), xz(new VariantConst(cur.getTiltFactorXZ())),
        yz(new VariantConst(cur.getTiltFactorYZ()));
    BoxResizeUpdater up(s->sysdef, vx, vy, vz, xy, xz, yz);
    up.setParams(scale_particles != 0);
    up.update(s->timestep);
    REF_CATCH(s, 1)
    return 0;
    }

void ref_get_box(void* h, float* L3)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    Scalar3 L = s->pdata->getGlobalBox().getL();
    L3[0] = L.x; L3[1] = L.y; L3[2] = L.z;
    }

//! the method's log quantity nvt_mtk_reservoir_energy (hoomd/md/TwoStepNVTMTK.cc:82-107)
int ref_get_nvt_reservoir_energy(void* h, float* out)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    bool mine = false;
    *out = s->nvt->getLogValue("nvt_mtk_reservoir_energy", s->timestep, mine);
    if (!mine) throw std::runtime_error("nvt_mtk_reservoir_energy is not provided by the method");
    REF_CATCH(s, 1)
    return 0;
    }

//! method.set_params(kT=...) / a temperature variant evaluated by the caller: setT on the active method
int ref_set_kT(void* h, float kT)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    std::shared_ptr<Variant> T(new VariantConst(kT));
    if (s->langevin) s->langevin->setT(T);
    if (s->nvt) s->nvt->setT(T);
    if (s->npt) s->npt->setT(T);
    REF_CATCH(s, 1)
    return 0;
    }

//! md.nlist.cell(dist_check=...): NeighborList::setEvery(every, dist_check) (hoomd/md/nlist.py:188-196)
int ref_set_dist_check(void* h, int on)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->nlist->setEvery(s->check_period, on != 0);
    REF_CATCH(s, 1)
    return 0;
    }

//! hoomd.update.sort (period 0 disables); the first sort happens at the first step run, like System::run
void ref_set_sorter(void* h, unsigned int period)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    s->sorter_period = period;
    s->sorter_next = s->timestep;
    if (period && !s->sorter)
        s->sorter = std::shared_ptr<SFCPackUpdater>(new SFCPackUpdater(s->sysdef));
    }

void ref_set_timestep(void* h, unsigned int ts)
    {
    static_cast<RefSystem*>(h)->timestep = ts;
    }

unsigned int ref_get_timestep(void* h)
    {
    return static_cast<RefSystem*>(h)->timestep;
    }

//! Only the neighbor list (and the cell list under it) at timestep ts
int ref_compute_nlist(void* h, unsigned int ts)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->nlist->compute(ts);
    REF_CATCH(s, 1)
    return 0;
    }

//! ForceCompute::compute(ts) of the LJ pair force (builds the nlist first if needed)
int ref_compute_forces(void* h, unsigned int ts)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->lj->compute(ts);
    if (s->bond) s->bond->compute(ts);
    REF_CATCH(s, 1)
    return 0;
    }

//! force / virial arrays of the bond ForceCompute after ref_compute_forces
void ref_get_bond_forces(void* h, float* force4, float* virial6)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int N = s->pdata->getN();
    ArrayHandle<Scalar4> f(s->bond->getForceArray(), access_location::host, access_mode::read);
    memcpy(force4, f.data, sizeof(Scalar4) * N);
    if (virial6)
        {
        unsigned int pitch = s->bond->getVirialArray().getPitch();
        ArrayHandle<Scalar> v(s->bond->getVirialArray(), access_location::host, access_mode::read);
        for (unsigned int k = 0; k < 6; k++)
            memcpy(virial6 + k * N, v.data + k * pitch, sizeof(Scalar) * N);
        }
    }

//! System::run(nsteps): prepRun at the current step, then updaters + integrator->update per step
/*! Follows hoomd/System.cc:484-599. Returns elapsed seconds of the step loop (excluding prepRun) in *elapsed */
int ref_run(void* h, unsigned int nsteps, double* elapsed)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->integrator->prepRun(s->timestep);
    auto t0 = std::chrono::steady_clock::now();
    for (unsigned int k = 0; k < nsteps; k++)
        {
        if (s->sorter && s->sorter_period && s->timestep == s->sorter_next)
            {
            s->sorter->update(s->timestep);
            s->sorter_next += s->sorter_period;
            }
        s->integrator->update(s->timestep);
        s->timestep++;
        }
    auto t1 = std::chrono::steady_clock::now();
    if (elapsed) *elapsed = std::chrono::duration<double>(t1 - t0).count();
    REF_CATCH(s, 1)
    return 0;
    }

//! One isolated integration half step on the current state (for kernel-level parity of step one / two)
int ref_step_one(void* h, unsigned int ts)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    if (s->langevin) s->langevin->integrateStepOne(ts); else s->nve->integrateStepOne(ts);
    REF_CATCH(s, 1)
    return 0;
    }

int ref_step_two(void* h, unsigned int ts, const float* net_force4)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
        {
        ArrayHandle<Scalar4> h_nf(s->pdata->getNetForce(), access_location::host, access_mode::readwrite);
        memcpy(h_nf.data, net_force4, sizeof(Scalar4) * s->pdata->getN());
        }
    if (s->langevin) s->langevin->integrateStepTwo(ts); else s->nve->integrateStepTwo(ts);
    REF_CATCH(s, 1)
    return 0;
    }

// ---------------------------------------------------------------- getters
unsigned int ref_get_N(void* h)
    {
    return static_cast<RefSystem*>(h)->pdata->getN();
    }

void ref_get_particles(void* h, float* pos4, float* vel4, float* accel3, int* image3, unsigned int* tag,
                       unsigned int* rtag)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int N = s->pdata->getN();
    if (pos4)
        {
        ArrayHandle<Scalar4> a(s->pdata->getPositions(), access_location::host, access_mode::read);
        memcpy(pos4, a.data, sizeof(Scalar4) * N);
        }
    if (vel4)
        {
        ArrayHandle<Scalar4> a(s->pdata->getVelocities(), access_location::host, access_mode::read);
        memcpy(vel4, a.data, sizeof(Scalar4) * N);
        }
    if (accel3)
        {
        ArrayHandle<Scalar3> a(s->pdata->getAccelerations(), access_location::host, access_mode::read);
        memcpy(accel3, a.data, sizeof(Scalar3) * N);
        }
    if (image3)
        {
        ArrayHandle<int3> a(s->pdata->getImages(), access_location::host, access_mode::read);
        memcpy(image3, a.data, sizeof(int3) * N);
        }
    if (tag)
        {
        ArrayHandle<unsigned int> a(s->pdata->getTags(), access_location::host, access_mode::read);
        memcpy(tag, a.data, sizeof(unsigned int) * N);
        }
    if (rtag)
        {
        ArrayHandle<unsigned int> a(s->pdata->getRTags(), access_location::host, access_mode::read);
        memcpy(rtag, a.data, sizeof(unsigned int) * N);
        }
    }

//! force (fx,fy,fz,pe) per particle and the 6 x pitch virial of the LJ force compute
unsigned int ref_get_virial_pitch(void* h)
    {
    return static_cast<RefSystem*>(h)->lj->getVirialArray().getPitch();
    }

void ref_get_forces(void* h, float* force4, float* virial6)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int N = s->pdata->getN();
    ArrayHandle<Scalar4> f(s->lj->getForceArray(), access_location::host, access_mode::read);
    memcpy(force4, f.data, sizeof(Scalar4) * N);
    if (virial6)
        {
        unsigned int pitch = s->lj->getVirialArray().getPitch();
        ArrayHandle<Scalar> v(s->lj->getVirialArray(), access_location::host, access_mode::read);
        for (unsigned int k = 0; k < 6; k++)
            memcpy(virial6 + k * N, v.data + k * pitch, sizeof(Scalar) * N);
        }
    }

void ref_get_net_force(void* h, float* force4)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    ArrayHandle<Scalar4> f(s->pdata->getNetForce(), access_location::host, access_mode::read);
    memcpy(force4, f.data, sizeof(Scalar4) * s->pdata->getN());
    }

//! Neighbor list allocation size (per-type Nmax is recovered from head-list differences by the caller)
unsigned int ref_get_nlist_size(void* h)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    return s->nlist->getNListArray().getNumElements();
    }

void ref_get_nlist(void* h, unsigned int* n_neigh, unsigned int* head, unsigned int* nlist, unsigned int nlist_len)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int N = s->pdata->getN();
    if (n_neigh)
        {
        ArrayHandle<unsigned int> a(s->nlist->getNNeighArray(), access_location::host, access_mode::read);
        memcpy(n_neigh, a.data, sizeof(unsigned int) * N);
        }
    if (head)
        {
        ArrayHandle<unsigned int> a(s->nlist->getHeadList(), access_location::host, access_mode::read);
        memcpy(head, a.data, sizeof(unsigned int) * N);
        }
    if (nlist)
        {
        ArrayHandle<unsigned int> a(s->nlist->getNListArray(), access_location::host, access_mode::read);
        unsigned int n = std::min(nlist_len, (unsigned int)s->nlist->getNListArray().getNumElements());
        memcpy(nlist, a.data, sizeof(unsigned int) * n);
        }
    }

void ref_get_nlist_stats(void* h, unsigned long long* out3)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    out3[0] = s->nlist->getNumUpdates();
    out3[1] = s->nlist->getSmallestRebuild();
    out3[2] = 0;
    }

//! Cell list geometry: dim[3], Nmax, n_adj
void ref_get_cell_dims(void* h, unsigned int* out5)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    uint3 dim = s->cl->getDim();
    out5[0] = dim.x; out5[1] = dim.y; out5[2] = dim.z;
    out5[3] = s->cl->getNmax();
    out5[4] = s->cl->getCellAdjIndexer().getW();
    }

void ref_get_cell_list(void* h, unsigned int* cell_size, float* xyzf4, unsigned int* cell_adj)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    unsigned int ncell = s->cl->getCellIndexer().getNumElements();
    if (cell_size)
        {
        ArrayHandle<unsigned int> a(s->cl->getCellSizeArray(), access_location::host, access_mode::read);
        memcpy(cell_size, a.data, sizeof(unsigned int) * ncell);
        }
    if (xyzf4)
        {
        ArrayHandle<Scalar4> a(s->cl->getXYZFArray(), access_location::host, access_mode::read);
        memcpy(xyzf4, a.data, sizeof(Scalar4) * s->cl->getCellListIndexer().getNumElements());
        }
    if (cell_adj)
        {
        ArrayHandle<unsigned int> a(s->cl->getCellAdjArray(), access_location::host, access_mode::read);
        memcpy(cell_adj, a.data, sizeof(unsigned int) * s->cl->getCellAdjIndexer().getNumElements());
        }
    }

//! CellList::setComputeTDB(true) (what NeighborListGPUBinned asks for, hoomd/md/NeighborListGPUBinned.cc:30-33) and a
//! recompute of the cell list at \a ts; read the geometry (ref_get_cell_dims) and the arrays afterwards
int ref_cell_compute_tdb(void* h, unsigned int ts)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->cl->setComputeTDB(true);
    s->cl->compute(ts);
    REF_CATCH(s, 1)
    return 0;
    }

//! getTDBArray(): (type, diameter, body, 0) per cell-list slot, [Nmax * ncell] Scalar4
void ref_get_cell_tdb(void* h, float* tdb4)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    ArrayHandle<Scalar4> a(s->cl->getTDBArray(), access_location::host, access_mode::read);
    memcpy(tdb4, a.data, sizeof(Scalar4) * s->cl->getCellListIndexer().getNumElements());
    }

//! ComputeThermo at timestep ts: out = {KE, PE, temperature, pressure, ndof}
int ref_get_thermo(void* h, unsigned int ts, double* out5)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    s->thermo->setNDOF(s->integrator->getNDOF(s->group_all));
    s->thermo->compute(ts);
    out5[0] = s->thermo->getKineticEnergy();
    out5[1] = s->thermo->getPotentialEnergy();
    out5[2] = s->thermo->getTemperature();
    out5[3] = s->virial ? s->thermo->getPressure() : 0.0;
    out5[4] = s->thermo->getNDOF();
    REF_CATCH(s, 1)
    return 0;
    }

//! Double-precision sum of the LJ per-particle energies (ForceCompute::calcEnergySum, hoomd/ForceCompute.cc:82-102)
double ref_get_lj_energy(void* h)
    {
    return static_cast<RefSystem*>(h)->lj->calcEnergySum();
    }

// ---------------------------------------------------------------- stand-alone pieces of the reference headers
//! hoomd/Saru.h: n draws of u32() and of s<float>(-1,1) from Saru(seed1, seed2, seed3)
void ref_saru(unsigned int seed1, unsigned int seed2, unsigned int seed3, unsigned int n, unsigned int* u32_out,
              float* f_out)
    {
    if (u32_out)
        {
        hoomd::detail::Saru saru(seed1, seed2, seed3);
        for (unsigned int i = 0; i < n; i++) u32_out[i] = saru.u32();
        }
    if (f_out)
        {
        hoomd::detail::Saru saru(seed1, seed2, seed3);
        for (unsigned int i = 0; i < n; i++) f_out[i] = saru.s<float>(-1, 1);
        }
    }

//! hoomd/SFCPackUpdater: run one sort now (whatever the period) -- particle order afterwards via ref_get_particles(tag)
int ref_sort_now(void* h)
    {
    RefSystem* s = static_cast<RefSystem*>(h);
    REF_TRY(s)
    if (!s->sorter) s->sorter = std::shared_ptr<SFCPackUpdater>(new SFCPackUpdater(s->sysdef));
    s->sorter->update(s->timestep);
    REF_CATCH(s, 1)
    return 0;
    }

} // extern "C"
