// stand-in forwarding header (test infrastructure, see sr_cv_standin.hpp)
#pragma once
#include "../sr_cv_standin.hpp"
