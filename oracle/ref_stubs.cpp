// ORACLE / TEST INFRASTRUCTURE ONLY.
// The reference's src/Utils/Utils.cpp pulls in the OIDN denoiser; the CPU kernels only need
// Utils::debugbreak() (src/Device/kernels/FullPathTracer.h:88), stubbed here as a no-op.
#include "Utils/Utils.h"
void Utils::debugbreak() {}
