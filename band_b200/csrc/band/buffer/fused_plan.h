// band/buffer/fused_plan.h -- the plan compiler between the operator classes
// and the CUDA shim (include/bandcu.h).
//
// The reference runs a processor as one materialised libyuv pass per operator
// (/root/reference/band/buffer/buffer_processor.cc:22-43 ->
// libyuv_image_operator.cc). Here the operator chain is LOWERED once per input
// signature into the fewest possible launches: Crop/Rotate/Flip become integer
// index maps, ColorSpaceConvert a per-fetch conversion, Resize the sampling
// stage and Normalize/DataTypeConvert the store stage of ONE kernel that writes
// straight into the tensor. A stage that cannot legally fold into the current
// launch (an index map after a Resize, a second conversion, ...) closes the
// segment into device scratch and starts the next one, so every chain the
// reference accepts still produces stage-by-stage identical values.
#ifndef BAND_BUFFER_FUSED_PLAN_H_
#define BAND_BUFFER_FUSED_PLAN_H_

#include <memory>
#include <string>
#include <vector>

#include "absl/status/status.h"
#include "absl/status/statusor.h"
#include "band/buffer/buffer.h"
#include "band/buffer/host_registry.h"
#include "band/buffer/scale_dispatch.h"
#include "bandcu.h"

namespace band {
class IBufferOperator;

namespace cuda {

// Shape + layout of a materialised image (chain input, scratch or chain output).
struct ImageLayout {
  BufferFormat format = BufferFormat::kRaw;
  DataType dtype = DataType::kUInt8;
  int width = 0, height = 0;
  size_t row_stride[3] = {0, 0, 0};
  size_t pixel_stride[3] = {0, 0, 0};
  int num_planes = 0;
  int orientation = 1;  // BufferOrientation: AutoOrient lowers differently per value, so it keys the plan
  static ImageLayout Of(const Buffer& b);
  bool operator==(const ImageLayout& o) const;
};

enum : int { kSlotInput = -1, kSlotOutput = -2 };  // otherwise: index into FusedPlan::scratch

struct Segment {
  enum class Kind { kPixel, kYuvMap, kYuvScale, kRgbToYuv, kElementwise };
  Kind kind = Kind::kPixel;
  int src_slot = kSlotInput;
  int dst_slot = kSlotOutput;
  // kPixel
  BandCuPlanDesc pixel{};
  BandCuPlan* plan = nullptr;
  // kYuvMap
  BandCuYuvMapDesc yuv{};
  // kRgbToYuv (ColorSpaceConvert RGB/RGBA -> NV12/NV21/YV12/YV21, libyuv_image_operator.cc:523-587,621-674)
  BandCuRgbToYuvDesc rgb2yuv{};
  // kYuvScale (Resize of a 4:2:0 buffer: I420Scale / NV12Scale, libyuv_image_operator.cc:376-435): one plan per
  // plane -- [0] luma, [1] U (or the interleaved UV plane), [2] V (planar only)
  BandCuPlan* plane_plan[3] = {nullptr, nullptr, nullptr};
  bool interleaved_uv = false;
  // kElementwise (Normalize / DataTypeConvert on an already materialised buffer)
  int in_dtype = 0, out_dtype = 0;
  long long count = 0;
  float mean = 0.f, std = 1.f;
};

// One frame's resolved device pointers for a slot.
struct SlotPointers {
  unsigned char* y = nullptr;  // plane 0
  unsigned char* u = nullptr;
  unsigned char* v = nullptr;
};

class FusedPlan {
 public:
  ~FusedPlan();
  // Runs n frames. inputs/outputs may live in host or device memory (all frames
  // of a call in the same space); host frames are staged through plan-owned
  // device memory. Synchronous: returns after the stream has drained.
  absl::Status Run(const Buffer* const* inputs, Buffer* const* outputs, int n, bool synchronize = true);
  // Launches the previous Run() again -- same frames, same outputs, no validation, no pointer resolution. Only legal when
  // CanRerun(): the previous Run was one pass over device-resident or in-place (pinned) memory. The caller
  // (BufferProcessor) establishes that the frame lists name the same views as before (Buffer::Epoch()).
  bool CanRerun() const { return rerunnable_; }
  absl::Status Rerun(bool synchronize);
  // Waits for everything a Run(..., false) enqueued (compute and copy-out streams).
  absl::Status Synchronize();
  // Records one Run() over DEVICE-resident (or pinned, zero-copy) frames into a CUDA graph and returns the executable
  // graph (bandcu_graph_launch / bandcu_graph_destroy). Replaying it costs one graph launch: no validation, no pointer
  // resolution, no per-segment launch calls. The frames and outputs must stay where they are while the graph is used,
  // and the graph reads plan-owned device memory (Normalize table, scratch images): whoever keeps the graph keeps a
  // shared_ptr to this plan (band::RecordedBatch). A batch that would be staged through copies is refused.
  absl::Status Record(const Buffer* const* inputs, Buffer* const* outputs, int n, void** graph_exec);
  // The stream the kernels are launched on (created on first use); for CUDA-event timing by harnesses.
  void* ComputeStream();
  size_t NumSegments() const { return segments_.size(); }
  bool IsSingleFastLaunch() const;
  const ImageLayout& input_layout() const { return input_; }
  const ImageLayout& output_layout() const { return output_; }
  bool writes_output() const { return writes_output_; }

 private:
  friend class ChainLowering;
  FusedPlan() = default;
  FusedPlan(const FusedPlan&) = delete;
  absl::Status EnsureStreams();
  absl::Status EnsureCapacity(int frames, size_t in_bytes, size_t out_bytes);
  absl::Status RunPass(const Buffer* const* inputs, Buffer* const* outputs, int m, bool in_host, bool out_host,
                       bool synchronize);
  absl::Status RunSegments(int first, int last, bool rerun = false);
  // the frame's planes are device-readable in place: pinned memory, or registered through the HostRegistry now
  bool HostFrameIsMapped(const Buffer& frame);
  void ReleaseHeldHost();
  SlotPointers Resolve(int slot, int frame) const;
  void* Event(size_t index);

  ImageLayout input_, output_;
  std::vector<ImageLayout> scratch_;
  std::vector<Segment> segments_;
  bool writes_output_ = true;   // false: chain ends in an in-place stage (output untouched, as in the reference)
  bool writes_input_ = false;   // an in-place stage targets the chain input
  // device memory owned by the plan
  int capacity_ = 0;                         // frames the scratch images hold
  std::vector<unsigned char*> scratch_mem_;  // per scratch image: capacity_ * scratch_bytes_
  std::vector<size_t> scratch_bytes_;
  unsigned char* in_stage_ = nullptr;   // host frames are uploaded here
  size_t in_stage_bytes_ = 0;
  unsigned char* out_stage_ = nullptr;  // results bound for host tensors land here first
  size_t out_stage_bytes_ = 0;
  // per-call frame pointers (chain input / output), device addresses
  std::vector<SlotPointers> in_ptrs_, out_ptrs_;
  std::vector<BandCuFrame> frames_;
  // compute stream + one copy stream per direction, so uploads, kernels and
  // downloads of consecutive chunks overlap
  void* stream_ = nullptr;
  void* copy_in_ = nullptr;
  void* copy_out_ = nullptr;
  std::vector<void*> events_;
  void* fence_a_ = nullptr;  // orders a new pass's uploads after the previous pass's kernels
  void* fence_b_ = nullptr;  // orders a new pass's kernels after the previous pass's downloads
  bool in_flight_ = false;   // an unsynchronised pass may still be running
  bool last_pass_staged_ = false;  // the last RunPass moved frames or results through staging copies
  bool rerunnable_ = false;        // the last Run() was a single pointer-only pass: Rerun() may replay it
  int last_frames_ = 0;
  // environment switches, read ONCE when the plan is created (never on the launch path)
  bool zero_copy_allowed_ = true;  // BAND_NO_ZERO_COPY
  int chunk_frames_ = 32;          // BAND_CHUNK_FRAMES: frames per upload/compute/download pipeline stage
  bool register_host_ = false;     // HostRegistry was on when the plan was created
  std::vector<HostRegistry::Entry*> held_host_;  // registered ranges the passes in flight read
};

// Walks an operator chain and emits segments. Operators call the stage
// methods from their Lower() hook.
class ChainLowering {
 public:
  explicit ChainLowering(const Buffer& input);
  ~ChainLowering();

  // index stages (dst pixel <- src pixel), reference semantics cited at each operator
  absl::Status Crop(int x0, int y0, int x1, int y1);
  absl::Status Rotate(int angle_ccw_deg);
  absl::Status Flip(bool horizontal, bool vertical);
  absl::Status ConvertColor(BufferFormat to);
  absl::Status Scale(int dst_width, int dst_height);
  absl::Status Normalize(float mean, float std, DataType out_dtype, bool inplace);
  // a stage that leaves its output untouched (DataTypeConvert between equal types,
  // common_operator.cc:169-175): the chain passes through unchanged
  absl::Status PassThrough();
  // lowers a nested operator (AutoConvert's sub-operators)
  absl::Status LowerNested(const IBufferOperator& op, const Buffer& input);

  absl::StatusOr<std::unique_ptr<FusedPlan>> Finish(const Buffer* output);

 private:
  struct View;
  absl::Status Flush();
  absl::Status EmitInto(int dst_slot, const ImageLayout& dst);
  absl::Status ApplyIndexMap(const BandCuAffine& luma, const BandCuAffine& chroma, int new_w, int new_h);
  void ResetView(int slot, const ImageLayout& layout);

  std::unique_ptr<FusedPlan> plan_;
  std::unique_ptr<View> view_;
  // the last stage wrote nothing the chain output should receive (in-place
  // Normalize or a skipped DataTypeConvert): the reference leaves `output` untouched
  bool terminal_noop_ = false;
};

// IBufferOperator::ProcessImpl for every operator: compile + run a one-stage chain. The compiled plan stays with the
// operator instance and is reused while the input and output layouts repeat.
absl::Status RunSingleOperator(IBufferOperator& op, const Buffer& input);

// Additive process-wide knobs of the GPU backing (the reference has no device notion).
absl::Status SetDevice(int device);

}  // namespace cuda
}  // namespace band
#endif  // BAND_BUFFER_FUSED_PLAN_H_
