// Programmatic dependent launch (griddepcontrol) for the few-frames schedule: a kernel launched with
// cudaLaunchAttributeProgrammaticStreamSerialization may start while the kernel before it on the stream still runs; its blocks
// do the part of their work that does not depend on that kernel (tap tables, box arithmetic, constant tables) and then call
// pdl_wait(), which returns once the previous kernel has completed and its writes are visible.  The hand-over gap between two
// dependent launches - most of a short kernel's cost when a call holds one frame - overlaps with that prologue.
// Both instructions are no-ops in a kernel that was launched the ordinary way.
#pragma once
#include <cuda_runtime.h>

namespace orbx {

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// lets the next kernel on the stream start its prologue (it still waits in its own pdl_wait for this grid to finish)
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

}  // namespace orbx
