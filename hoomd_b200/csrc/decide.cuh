// NeighborList::needsUpdating on the device, shared by the single-GPU step kernels and the multi-GPU step barrier.
#pragma once
#include "kernels.h"

namespace b200
{
// NeighborList::needsUpdating(ts+1) for the force evaluation that follows (update(ts) computes forces at ts+1).
// IN_MEMORY: `st` is the device-resident block (`moved` was set by other CTAs: re-read it); otherwise a register copy.
template<bool IN_MEMORY = true> __device__ __forceinline__ void decide_body(StepState* st)
    {
    const unsigned int ts1 = st->timestep + 1u;
    bool result;
    if (st->force_update)
        {
        result = true;
        st->force_update = 0u;
        st->num_forced += 1u;
        st->last_updated = ts1;
        }
    else if (ts1 < st->last_updated + st->every) // !shouldCheckDistance
        result = false;
    else
        {
        // distance checks off: the check period decides (every == 0: always); otherwise the distance check of the step
        const bool period_hit = st->no_dist_check && (st->every == 0u || (st->every > 1u && ts1 == st->last_updated + st->every));
        result = st->r_buff_tiny || period_hit || (IN_MEMORY ? *((volatile unsigned int*)&st->moved) : st->moved);
        if (result)
            {
            st->last_updated = ts1;
            st->num_updates += 1u;
            }
        }
    st->moved = 0u;
    if (result) st->box_scaled = 0u; // the build that follows resets m_last_L: lambda is 1 from the next check on
    st->need_rebuild = result ? 1u : 0u;
    st->ts_step_two = st->timestep;
    st->timestep = ts1;
    }

} // namespace b200
