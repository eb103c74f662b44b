// Test-infrastructure shim: lets the reference's hot-path sources compile without DuckDB.
// Everything lives in lmd_duckdb_shim.hpp; this file only forwards to it.
#pragma once
#include "duckdb/lmd_duckdb_shim.hpp"
