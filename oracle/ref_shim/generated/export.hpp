// Build shim for oracle/_ref: CMake normally generates this header.
#pragma once
#define EAR_EXPORT
