/* TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * Force-included (-include) into every translation unit of the *unmodified* reference sources
 * when oracle/Makefile builds oracle/_ref/librectify_canon.so.  The reference seeds its RANSAC
 * generator from std::random_device (reference src/estimator.h:35), which makes the reference
 * non-deterministic.  This shim swaps the type for a functor returning a process-global seed so
 * that the compiled reference becomes a deterministic checker.  Nothing else is altered.
 */
#pragma once
#include <random>
extern "C" unsigned oracle_ref_seed_value;
namespace std {
struct oracle_fixed_random_device {
    typedef unsigned result_type;
    result_type operator()() { return oracle_ref_seed_value; }
};
}
#define random_device oracle_fixed_random_device
