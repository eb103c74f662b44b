// Build-time stand-in (see duckdb/lmd_duckdb_sql_shim.hpp): forwards to it so that DuckDB's include paths resolve.
#pragma once
#include "duckdb/lmd_duckdb_sql_shim.hpp"
// DuckDB's C API header (duckdb.h), which the real umbrella header pulls in, declares idx_t at global scope; the
// reference's db/LmDiskannIndex.hpp relies on that (it names `idx_t` unqualified inside namespace diskann::db).
typedef uint64_t idx_t;
