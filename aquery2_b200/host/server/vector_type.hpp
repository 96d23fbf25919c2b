// drop-in for the reference header server/vector_type.hpp: everything lives in ../aqgpu_host.hpp
#pragma once
#include "../aqgpu_host.hpp"
