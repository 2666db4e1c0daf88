// SimFlags bitmask of the reference (src/sim_flags.hpp:7-13, sim_flags.inl).
#pragma once

#include <cstdint>

namespace GPUHideSeek {

enum class SimFlags : uint32_t {
    Default = 0,
    UseFixedWorld = 1 << 0,
    IgnoreEpisodeLength = 1 << 1,
    RandomFlipTeams = 1 << 2,
    ZeroAgentVelocity = 1 << 3,
};

inline SimFlags &operator|=(SimFlags &a, SimFlags b) { a = SimFlags(uint32_t(a) | uint32_t(b)); return a; }
inline SimFlags operator|(SimFlags a, SimFlags b) { a |= b; return a; }
inline SimFlags &operator&=(SimFlags &a, SimFlags b) { a = SimFlags(uint32_t(a) & uint32_t(b)); return a; }
inline SimFlags operator&(SimFlags a, SimFlags b) { a &= b; return a; }

}  // namespace GPUHideSeek
