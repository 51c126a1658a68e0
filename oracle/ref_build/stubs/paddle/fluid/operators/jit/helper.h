// Stub standing in for the reference's jit/helper.h while compiling
// paddle/fluid/operators/jit/refer/refer.h from /root/reference for oracle/_ref.
// refer.h uses nothing from the real helper.h (kernel pools, places); the real one
// drags in the whole framework header closure, which is not buildable here.
#pragma once
