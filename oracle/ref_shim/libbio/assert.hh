/*
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * Boost-free stand-in for the reference's libbio/assert.hh, written for this repo (the real header,
 * lib/libbio/include/libbio/assert.hh:9-10, needs boost/stacktrace.hpp which this image lacks).
 * It provides only the macros the hot-path sources use:
 *   rewrite_cigar.hh:102-103,150; msa_index.hh:23,223; cigar.hh:146; indel_run_checker.cc:97.
 * Failures print a message and abort(), like libbio_fail in the reference (or throw, with PV3_SHIM_FAIL_THROWS).
 */
#ifndef PV3_REF_SHIM_LIBBIO_ASSERT_HH
#define PV3_REF_SHIM_LIBBIO_ASSERT_HH

#include <cstdlib>
#include <iostream>

namespace libbio { namespace shim {
	struct failure {};

	template <typename... t_args>
	[[noreturn]] void fail(char const *file, int line, t_args const &... args)
	{
		std::cerr << file << ':' << line << ": ";
		((std::cerr << args), ...);
		std::cerr << std::endl;
		std::abort();
	}

	// For drivers that report the reference's aborts as per-record flags (oracle/ref_driver_mapq.cc). A separate name: the
	// two behaviours live in one shared object.
	template <typename... t_args>
	[[noreturn]] void fail_throwing(char const *, int, t_args const &...)
	{
		throw failure{};
	}
}}

#ifdef PV3_SHIM_FAIL_THROWS
#	define libbio_fail(...)			::libbio::shim::fail_throwing(__FILE__, __LINE__, __VA_ARGS__)
#else
#define libbio_fail(...)				::libbio::shim::fail(__FILE__, __LINE__, __VA_ARGS__)
#endif
#define libbio_always_assert(X)			do { if (!(X)) libbio_fail("assertion failed: ", #X); } while (false)
#define libbio_always_assert_lt(A, B)	libbio_always_assert((A) < (B))
#define libbio_always_assert_lte(A, B)	libbio_always_assert((A) <= (B))
#define libbio_always_assert_eq(A, B)	libbio_always_assert((A) == (B))
#define libbio_always_assert_neq(A, B)	libbio_always_assert((A) != (B))
#define libbio_always_assert_lte_(A, B)	libbio_always_assert((A) <= (B))

#ifdef LIBBIO_NDEBUG
#	define libbio_assert(X)			do {} while (false)
#	define libbio_assert_lt(A, B)		do {} while (false)
#	define libbio_assert_lte(A, B)		do {} while (false)
#	define libbio_assert_eq(A, B)		do {} while (false)
#	define libbio_assert_neq(A, B)		do {} while (false)
#else
#	define libbio_assert(X)			libbio_always_assert(X)
#	define libbio_assert_lt(A, B)		libbio_always_assert_lt(A, B)
#	define libbio_assert_lte(A, B)		libbio_always_assert_lte(A, B)
#	define libbio_assert_eq(A, B)		libbio_always_assert_eq(A, B)
#	define libbio_assert_neq(A, B)		libbio_always_assert_neq(A, B)
#endif

#endif
