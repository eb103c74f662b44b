// Build-time stand-in (see duckdb/lmd_duckdb_sql_shim.hpp): forwards to it so that DuckDB's include paths resolve.
#pragma once
#include "duckdb/lmd_duckdb_sql_shim.hpp"
