// BUILD-TIME STAND-IN, not product code: with real DuckDB v1.2.x headers on the include path this directory is
// simply left out. Used by oracle/Makefile (the reference's own sources) and by duckdb-lm-diskann_b200/host/Makefile
// (the host adapter written against the reference's interface headers).
//
// Minimal stand-in for the handful of DuckDB v1.2.x declarations that the reference's hot-path
// sources (core/Searcher.cpp, core/index_config.cpp, core/GraphNode.cpp, core/distance.hpp,
// common/ann.hpp, common/duckdb_types.hpp) name. DuckDB itself is absent from this image
// (SURVEY.md §0.4), so those sources are compiled against this shim.
// Nothing here is copied from DuckDB; each item is the smallest declaration with the same meaning.
#pragma once

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <mutex>
#include <random>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace duckdb {

using idx_t = uint64_t;
using row_t = int64_t;
using block_id_t = int64_t;
using data_t = uint8_t;
using data_ptr_t = data_t *;
using const_data_ptr_t = const data_t *;
using std::string;
using std::vector;
using std::unordered_map;
using std::mutex;
using std::atomic;
using std::unique_ptr;
using std::shared_ptr;

struct DConstants {
	static constexpr const idx_t INVALID_INDEX = idx_t(-1);
};

template <class T>
struct NumericLimits {
	static constexpr T Minimum() {
		return std::numeric_limits<T>::lowest();
	}
	static constexpr T Maximum() {
		return std::numeric_limits<T>::max();
	}
};

// DuckDB's IndexPointer is a 64-bit handle; 0 is what a default-constructed pointer holds and what
// the reference treats as "invalid" (Searcher.cpp:75).
class IndexPointer {
public:
	IndexPointer() : data(0) {
	}
	explicit IndexPointer(idx_t d) : data(d) {
	}
	idx_t Get() const {
		return data;
	}
	void Set(idx_t d) {
		data = d;
	}
	void Clear() {
		data = 0;
	}
	bool operator==(const IndexPointer &o) const {
		return data == o.data;
	}
	bool operator!=(const IndexPointer &o) const {
		return data != o.data;
	}

private:
	idx_t data;
};

enum class ExceptionType : uint8_t {
	INVALID = 0,
	OUT_OF_RANGE = 1,
	NOT_IMPLEMENTED = 11,
	IO = 18,
	INTERNAL = 31,
	INVALID_INPUT = 32,
	BINDER = 14
};

class Exception : public std::runtime_error {
public:
	Exception(ExceptionType t, const string &msg) : std::runtime_error(msg), type(t) {
	}
	ExceptionType type;
};
class InvalidInputException : public Exception {
public:
	explicit InvalidInputException(const string &msg) : Exception(ExceptionType::INVALID_INPUT, msg) {
	}
};
class InternalException : public Exception {
public:
	explicit InternalException(const string &msg) : Exception(ExceptionType::INTERNAL, msg) {
	}
};
class NotImplementedException : public Exception {
public:
	explicit NotImplementedException(const string &msg) : Exception(ExceptionType::NOT_IMPLEMENTED, msg) {
	}
};
class IOException : public Exception {
public:
	explicit IOException(const string &msg) : Exception(ExceptionType::IO, msg) {
	}
};
class BinderException : public Exception {
public:
	explicit BinderException(const string &msg) : Exception(ExceptionType::BINDER, msg) {
	}
};

class RandomEngine {
public:
	explicit RandomEngine(int64_t seed = 42) : gen(static_cast<uint32_t>(seed)) {
	}
	double NextRandom() {
		return std::uniform_real_distribution<double>(0.0, 1.0)(gen);
	}
	uint32_t NextRandomInteger() {
		return gen();
	}

private:
	std::mt19937 gen;
};

struct Printer {
	static void Print(const string &s) {
		std::fprintf(stderr, "%s\n", s.c_str());
	}
};

template <class T, T A = 8>
static inline constexpr T AlignValue(T n) {
	return ((n + (A - 1)) / A) * A;
}

struct Storage {
	static constexpr const idx_t SECTOR_SIZE = 4096;
};

struct StringUtil {
	static string Format(const char *fmt, ...) {
		char buf[1024];
		va_list ap;
		va_start(ap, fmt);
		std::vsnprintf(buf, sizeof(buf), fmt, ap);
		va_end(ap);
		return string(buf);
	}
};

} // namespace duckdb
