#pragma once
#include <mitsuba/mitsuba.h>  // mock: one header holds everything
