// Build shim for the oracle only: GPFunctions.cpp includes <fftw3.h> but calls nothing from it on this path.
#pragma once
