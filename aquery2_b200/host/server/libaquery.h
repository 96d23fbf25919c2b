// drop-in for the reference header server/libaquery.h: everything lives in ../aqgpu_host.hpp
#pragma once
#include "../aqgpu_host.hpp"
