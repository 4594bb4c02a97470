// Shared bits of the oracle/_ref harness. TEST INFRASTRUCTURE ONLY.
#pragma once
#include <cstdlib>
#include <new>
// ScalarLoss::Init reads nt_ before assigning it (/root/reference/learn/linear/loss.h:36-37).
// Zero-filled heap memory turns that read into num_threads(0) == "unspecified" for libgomp
// instead of garbage. Hidden visibility keeps the replacement inside this .so.
#define REF_HIDDEN __attribute__((visibility("hidden")))
#define REF_API extern "C" __attribute__((visibility("default")))
