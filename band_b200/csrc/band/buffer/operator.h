// band/buffer/operator.h -- IBufferOperator with the reference's protocol
// (/root/reference/band/buffer/operator.h:31-78, operator.cc:52-97):
//   Process = ValidateInput -> ValidateOrCreateOutput -> ProcessImpl,
// output caching, SetOutput. What changes underneath: an operator does no pixel
// work of its own. It DESCRIBES its stage to the plan compiler (Lower), and the
// compiled plan runs as fused CUDA launches. ProcessImpl of every operator is
// "compile and run the one-operator chain".
#ifndef BAND_BUFFER_OPERATOR_H_
#define BAND_BUFFER_OPERATOR_H_

#include <memory>

#include "absl/status/status.h"
#include "band/buffer/buffer.h"

namespace band {
class IBufferOperator;
namespace cuda {
class ChainLowering;
class FusedPlan;
absl::Status RunSingleOperator(IBufferOperator& op, const Buffer& input);
}

class IBufferOperator {
 public:
  IBufferOperator() = default;
  virtual ~IBufferOperator() = default;

  // Copies carry the parameters only: a copy never shares or duplicates an
  // output buffer (the reference shallow-copies an owning Buffer here,
  // operator.cc:25-36).
  IBufferOperator(const IBufferOperator& rhs);
  IBufferOperator(IBufferOperator&& rhs);
  IBufferOperator& operator=(const IBufferOperator& rhs);

  absl::Status Process(const Buffer& input);
  virtual void SetOutput(Buffer* output);
  Buffer* GetOutput();
  const Buffer* GetOutput() const;

  enum class Type {
    kImage,
    kCommon,
  };
  virtual Type GetOpType() const = 0;
  virtual absl::Status ValidateInput(const Buffer& input) const;
  absl::Status ValidateOrCreateOutput(const Buffer& input);
  // called only when the output buffer is not null
  virtual absl::Status ValidateOutput(const Buffer& input) const = 0;
  virtual absl::Status CreateOutput(const Buffer& input) = 0;

 protected:
  friend class ImageProcessorBuilder;
  friend class BufferProcessor;
  friend class cuda::ChainLowering;
  friend absl::Status cuda::RunSingleOperator(IBufferOperator& op, const Buffer& input);

  virtual absl::Status ProcessImpl(const Buffer& input);
  virtual IBufferOperator* Clone() const = 0;

  // Additive: append this operator's stage to a fused plan. `input` is the
  // (possibly virtual) buffer the stage consumes; GetOutput() is what it yields.
  virtual absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const = 0;

  // Intermediates created by CreateOutput: host memory for a stand-alone
  // operator, shape-only inside a processor (the fused plan never materialises them).
  MemorySpace OutputSpace() const { return virtual_outputs_ ? MemorySpace::kVirtual : MemorySpace::kHost; }
  void OwnOutput(Buffer* created);
  // nested operators (AutoConvert's) create the same kind of intermediates as their parent
  static void InheritOutputSpace(const IBufferOperator& parent, IBufferOperator& child) {
    child.virtual_outputs_ = parent.virtual_outputs_;
  }

  bool output_assigned_ = false;
  bool virtual_outputs_ = false;
  Buffer* output_ = nullptr;              // assigned (not owned) or == owned_output_.get()
  std::unique_ptr<Buffer> owned_output_;  // created by CreateOutput
  // stand-alone Process(): the compiled one-operator plan, kept while the layouts repeat (never copied with the operator)
  std::shared_ptr<cuda::FusedPlan> single_plan_;
};
}  // namespace band

#endif  // BAND_BUFFER_OPERATOR_H_
