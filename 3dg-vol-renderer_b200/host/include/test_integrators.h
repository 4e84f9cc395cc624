// test_integrators.h -- kept so that user code including the reference header name compiles;
// the classes live in march_integrators.h (pulled in by integrator.h).
#pragma once
#include "integrator.h"
