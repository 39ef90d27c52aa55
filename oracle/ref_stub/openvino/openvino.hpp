// Test infrastructure only (oracle/_ref build): a do-nothing stand-in for <openvino/openvino.hpp>
// so that the reference's reranker.cpp / post_processor.cpp / parse_inputs.cpp can be compiled
// IN PLACE from /root/reference without OpenVINO (absent in this image). Only the type names
// that includes/inference/fast_model.hpp mentions are declared; nothing here is ever executed.
#pragma once
#include <unordered_set>
#include <iomanip>
namespace ov {
struct Node {};
struct Core {};
struct CompiledModel {};
template <class T> struct Output {};
struct InferRequest {};
struct Tensor {};
}  // namespace ov
