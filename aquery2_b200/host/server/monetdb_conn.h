// drop-in for the reference header server/monetdb_conn.h: everything lives in ../aqgpu_host.hpp
#pragma once
#include "../aqgpu_host.hpp"
