// Replacement for dlib-dnn-pimpl-wrapper/NetPimpl.cpp: the SAME public classes (NetPimpl.h is used unchanged),
// implemented over the C ABI of libannonet_b200.so instead of dlib's dnn/cuda code.  Build it in place of the
// reference's NetPimpl.cpp; dlib is needed for its header-only types only (matrix,
// resizable_tensor, rgb_pixel, sgd, count_steps_without_decrease) and is compiled WITHOUT DLIB_USE_CUDA.
//
// What maps to what:
//   TrainingNet::Initialize/SetClassCount/SetNetWidth  -> deferred; ab_net_init_random at the first StartTraining
//                                                        (dlib allocates parameters at the first forward pass too)
//   TrainingNet::StartTraining                         -> ab_train_step_async   (depth-1 asynchrony, trainer.h:87,1252)
//   TrainingNet::GetLearningRate / GetRuntimeNet       -> ab_train_sync (+ plateau detection, trainer.h:835-864)
//   RuntimeNet::Forward / operator() / Process         -> ab_forward / ab_process
//   Serialize / Deserialize                            -> ab_net_save_dlib / ab_net_load_dlib_prefix (dlib streams)
//   errors                                             -> ab_status mapped back to the reference's exception types
#include "NetPimpl.h"

#include "annonet_b200.h"

#include <deque>
#include <fstream>
#include <iterator>
#include <sstream>

namespace NetPimpl {

namespace {

int env_int(const char* name, int fallback)
{
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : fallback;
}

// ab_status -> the exception type the reference would have thrown (include/annonet_b200.h)
void check(ab_status st)
{
    if (st == AB_OK) return;
    const std::string what = ab_last_error();
    switch (st) {
    case AB_ERR_SERIALIZATION: throw dlib::serialization_error(what);
    case AB_ERR_CUDA: throw dlib::cuda_error(what);
    case AB_ERR_TOO_LARGE: throw std::runtime_error(what);
    case AB_ERR_ARGUMENT: throw dlib::fatal_error(dlib::EBROKEN_ASSERT, what);
    default: throw dlib::error(what);
    }
}

ab_net* new_net()
{
    ab_net* n = nullptr;
    // AB_DEVICE / AB_PRECISION: deployment knobs of the engine (0 = fp32 parity path, 1 = bf16 tensor-core path)
    check(ab_net_create(env_int("AB_DEVICE", 0), env_int("AB_PRECISION", AB_PRECISION_F16), &n));
    return n;
}

std::string read_rest(std::istream& in)
{
    return std::string(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
}

// deserialize(net, in) leaves the stream right behind the net: read ahead, then seek back to the first unused byte
void load_from_stream(ab_net* net, std::istream& in)
{
    const std::istream::pos_type start = in.tellg();
    const std::string bytes = read_rest(in);
    size_t used = 0;
    check(ab_net_load_dlib_prefix(net, bytes.data(), bytes.size(), &used));
    if (start != std::istream::pos_type(-1)) {
        in.clear();
        in.seekg(start + std::istream::off_type(used));
    }
}

void save_to_stream(ab_net* net, int as_runtime, std::ostream& out)
{
    void* bytes = nullptr;
    size_t n = 0;
    check(ab_net_save_dlib(net, as_runtime, &bytes, &n));
    out.write(static_cast<const char*>(bytes), (std::streamsize)n);
    ab_free(bytes);
}

}  // namespace

struct TrainingNet::Impl {
    ab_net* net = nullptr;
    bool allocated = false;          // parameters exist (first StartTraining or Deserialize)
    bool step_pending = false;
    // configuration collected before the parameters are allocated
    int class_count = 2;
    double width_scaler = 1.0;
    int min_filter_count = 1;
    double learning_rate = 1e-2;     // dnn_trainer default (trainer.h:78)
    double weight_decay = 0.001, momentum = 0.9;
    double shrink = 0.1;
    unsigned long iter_without_progress = 2000, dump_amount = 400, bn_window = 100;
    bool verbose = false;
    // plateau detection state (trainer.h:612-623,835-864)
    std::deque<double> previous_loss_values;
    unsigned long gradient_check_budget = 0, steps = 0;
    // staging
    std::vector<uint8_t> images;
    std::vector<ab_wlabel> labels;

    ~Impl() { if (net) ab_net_destroy(net); }

    void push_hyper() const
    {
        check(ab_set_hyper(net, AB_HYPER_LEARNING_RATE, learning_rate));
        check(ab_set_hyper(net, AB_HYPER_WEIGHT_DECAY, weight_decay));
        check(ab_set_hyper(net, AB_HYPER_MOMENTUM, momentum));
        check(ab_set_hyper(net, AB_HYPER_BN_STATS_WINDOW, (double)bn_window));
    }

    // the bookkeeping dnn_trainer's thread does after each step (record_loss + learning-rate shrink on a plateau)
    void finish_step()
    {
        if (!step_pending) return;
        double loss = 0;
        check(ab_train_sync(net, &loss));
        step_pending = false;
        ++steps;
        gradient_check_budget += 200;
        previous_loss_values.push_back(loss);
        while (previous_loss_values.size() > iter_without_progress) previous_loss_values.pop_front();
        if (verbose && steps % 20 == 0)
            std::cout << "step#: " << steps << "  learning rate: " << learning_rate << "  train loss: " << loss << std::endl;
        if (gradient_check_budget > iter_without_progress && shrink != 1) {
            gradient_check_budget = 0;
            if (dlib::count_steps_without_decrease(previous_loss_values) >= iter_without_progress &&
                dlib::count_steps_without_decrease_robust(previous_loss_values) >= iter_without_progress) {
                learning_rate *= shrink;
                check(ab_set_hyper(net, AB_HYPER_LEARNING_RATE, learning_rate));
                for (unsigned long cnt = 0; cnt < dump_amount + iter_without_progress / 10 && !previous_loss_values.empty(); ++cnt)
                    previous_loss_values.pop_front();
            }
        }
    }
};

struct RuntimeNet::Impl {
    ab_net* net = nullptr;
    dlib::resizable_tensor output;   // Forward()/GetOutput() hand out a host-readable tensor (annonet_infer.cpp:100)
    ~Impl() { if (net) ab_net_destroy(net); }
};

// ---- TrainingNet ---------------------------------------------------------------------------------
TrainingNet::TrainingNet() { pimpl = new Impl(); }
TrainingNet::~TrainingNet() { delete pimpl; }

void TrainingNet::Initialize(const solver_type& solver, const std::vector<int> extraDevices, std::shared_ptr<ThreadPools>)
{
    if (!extraDevices.empty())
        throw dlib::error("annonet_b200: one process drives one GPU; run data-parallel training as one process per GPU "
                          "(torchrun / mpirun) and call ab_comm_init -- see INTEGRATION.md");
    if (pimpl->net) { pimpl->finish_step(); ab_net_destroy(pimpl->net); }
    const unsigned long window = pimpl->bn_window;
    *pimpl = Impl();
    pimpl->bn_window = window;
    pimpl->net = new_net();
    pimpl->weight_decay = solver.get_weight_decay();
    pimpl->momentum = solver.get_momentum();
}

void TrainingNet::SetClassCount(unsigned short classCount)
{
    if (classCount >= dlib::loss_multiclass_log_per_pixel_::label_to_ignore)
        throw dlib::fatal_error(dlib::EBROKEN_ASSERT, "classCount < label_to_ignore");
    pimpl->class_count = classCount;
}
void TrainingNet::SetLearningRate(double learningRate)
{
    pimpl->finish_step();
    pimpl->learning_rate = learningRate;
    pimpl->previous_loss_values.clear();   // dnn_trainer::set_learning_rate resets the progress counters (trainer.h:455-463)
    if (pimpl->allocated) check(ab_set_hyper(pimpl->net, AB_HYPER_LEARNING_RATE, learningRate));
}
void TrainingNet::SetIterationsWithoutProgressThreshold(unsigned long threshold) { pimpl->iter_without_progress = threshold; }
void TrainingNet::SetPreviousLossValuesDumpAmount(unsigned long dump_amount) { pimpl->dump_amount = dump_amount; }
void TrainingNet::SetAllBatchNormalizationRunningStatsWindowSizes(unsigned long window_size)
{
    if (window_size == 0) throw dlib::fatal_error(dlib::EBROKEN_ASSERT, "The batch normalization running stats window size can't be 0.");
    pimpl->bn_window = window_size;
    if (pimpl->allocated) check(ab_set_hyper(pimpl->net, AB_HYPER_BN_STATS_WINDOW, (double)window_size));
}
void TrainingNet::SetLearningRateShrinkFactor(double f) { pimpl->shrink = f; }
void TrainingNet::SetSynchronizationFile(const std::string&, std::chrono::seconds)
{
    // dnn_trainer's own state file (trainer.h:931-1113) is a "next" row: the CLIs also save the net itself
    // (annonet_train_main.cpp:557-565), which this build supports through Serialize.
}
void TrainingNet::SetNetWidth(double scaler, int minFilterCount)
{
    pimpl->width_scaler = scaler;
    pimpl->min_filter_count = minFilterCount;
}
void TrainingNet::BeVerbose() { pimpl->verbose = true; }

int TrainingNet::GetRequiredInputDimension() { return ab_required_input_dimension(); }

void TrainingNet::StartTraining(const std::vector<input_type>& inputs, const std::vector<training_label_type>& training_labels)
{
    if (!pimpl->net) Initialize();
    if (inputs.empty() || inputs.size() != training_labels.size())
        throw dlib::fatal_error(dlib::EBROKEN_ASSERT, "inputs and training_labels must be non-empty and of equal size");
    pimpl->finish_step();   // rendezvous: the previous step has been taken off the pipe (trainer.h:1252)
    if (!pimpl->allocated) {
        check(ab_net_init_random(pimpl->net, pimpl->class_count, pimpl->width_scaler, pimpl->min_filter_count, (uint64_t)std::rand()));
        pimpl->allocated = true;
    }
    pimpl->push_hyper();
    const long h = inputs[0].nr(), w = inputs[0].nc();
    const size_t n = inputs.size(), px = (size_t)h * w;
    pimpl->images.resize(n * px * 3);
    pimpl->labels.resize(n * px);
    static_assert(sizeof(dlib::rgb_pixel) == 3, "rgb_pixel is 3 interleaved bytes");
    static_assert(sizeof(training_label_type::type) == sizeof(ab_wlabel), "weighted_label<uint16_t> is {u16, f32}");
    for (size_t i = 0; i < n; ++i) {
        if (inputs[i].nr() != h || inputs[i].nc() != w || training_labels[i].nr() != h || training_labels[i].nc() != w)
            throw dlib::fatal_error(dlib::EBROKEN_ASSERT, "all crops and label images of a mini-batch must have the same size");
        std::memcpy(pimpl->images.data() + i * px * 3, &inputs[i](0, 0), px * 3);
        for (size_t p = 0; p < px; ++p) {
            const auto& l = training_labels[i](p / w, p % w);
            pimpl->labels[i * px + p].label = l.label;
            pimpl->labels[i * px + p].weight = l.weight;
        }
    }
    check(ab_train_step_async(pimpl->net, pimpl->images.data(), pimpl->labels.data(), (int)n, (int)h, (int)w));
    pimpl->step_pending = true;
}

double TrainingNet::GetLearningRate() const
{
    pimpl->finish_step();
    return pimpl->learning_rate;
}

RuntimeNet TrainingNet::GetRuntimeNet() const
{
    RuntimeNet r;
    r = *this;
    return r;
}

void TrainingNet::Serialize(std::ostream& out) const
{
    pimpl->finish_step();
    if (!pimpl->allocated) throw dlib::serialization_error("annonet_b200: the net's parameters are not allocated yet (no training step has run)");
    save_to_stream(pimpl->net, 0, out);
}
void TrainingNet::Deserialize(std::istream& in)
{
    if (!pimpl->net) pimpl->net = new_net();
    pimpl->finish_step();
    load_from_stream(pimpl->net, in);
    pimpl->allocated = true;
}
void TrainingNet::Serialize(const std::string& filename) const
{
    std::ofstream out(filename, std::ios::binary);
    Serialize(out);
}
void TrainingNet::Deserialize(const std::string& filename)
{
    std::ifstream in(filename, std::ios::binary);
    Deserialize(in);
}

// ---- RuntimeNet ----------------------------------------------------------------------------------
RuntimeNet::RuntimeNet() { pimpl = new Impl(); }
RuntimeNet::~RuntimeNet() { delete pimpl; }
RuntimeNet::RuntimeNet(const RuntimeNet& that)
{
    pimpl = new Impl();
    operator=(that);
}
RuntimeNet& RuntimeNet::operator=(const RuntimeNet& that)
{
    if (this == &that) return *this;
    if (pimpl->net) { ab_net_destroy(pimpl->net); pimpl->net = nullptr; }
    if (that.pimpl->net) check(ab_net_clone(that.pimpl->net, &pimpl->net));
    return *this;
}
RuntimeNet& RuntimeNet::operator=(const TrainingNet& trainingNet)
{
    TrainingNet::Impl* t = trainingNet.pimpl;
    t->finish_step();   // may block (get_net)
    if (!t->allocated) throw dlib::error("annonet_b200: the training net has no parameters yet");
    // bn_ -> affine_ with the running statistics folded in (layers.h:2144-2163), through the serialized form
    void* bytes = nullptr;
    size_t n = 0;
    check(ab_net_save_dlib(t->net, 1, &bytes, &n));
    if (!pimpl->net) pimpl->net = new_net();
    const ab_status st = ab_net_load_dlib(pimpl->net, bytes, n);
    ab_free(bytes);
    check(st);
    return *this;
}

static void require_loaded(ab_net* net)
{
    if (!net) throw dlib::error("annonet_b200: the runtime net holds no parameters (Deserialize it or assign a TrainingNet first)");
}

output_type RuntimeNet::operator()(const input_type& input, const std::vector<double>& gainFactors) const { return Process(input, gainFactors); }

output_type RuntimeNet::Process(const input_type& input, const std::vector<double>& gainFactors) const
{
    require_loaded(pimpl->net);
    output_type labels(input.nr(), input.nc());
    check(ab_process(pimpl->net, reinterpret_cast<const uint8_t*>(&input(0, 0)), (int)input.nr(), (int)input.nc(),
                     gainFactors.empty() ? nullptr : gainFactors.data(), (int)gainFactors.size(), &labels(0, 0)));
    return labels;
}

const dlib::tensor& RuntimeNet::GetOutput() const { return pimpl->output; }

int RuntimeNet::GetRecommendedInputDimension(int minimumInputDimension)
{
    int out = 0;
    check(ab_recommended_input_dimension(minimumInputDimension, &out));
    return out;
}

const dlib::tensor& RuntimeNet::Forward(const input_type& input) const
{
    require_loaded(pimpl->net);
    ab_net_info_t info;
    check(ab_net_info(pimpl->net, &info));
    pimpl->output.set_size(1, info.class_count, input.nr(), input.nc());
    check(ab_forward(pimpl->net, reinterpret_cast<const uint8_t*>(&input(0, 0)), (int)input.nr(), (int)input.nc(), pimpl->output.host_write_only()));
    return pimpl->output;
}

void RuntimeNet::Serialize(std::ostream& out) const
{
    require_loaded(pimpl->net);
    save_to_stream(pimpl->net, 0, out);
}
void RuntimeNet::Deserialize(std::istream& in)
{
    if (!pimpl->net) pimpl->net = new_net();
    load_from_stream(pimpl->net, in);
}
void RuntimeNet::Serialize(const std::string& filename) const
{
    std::ofstream out(filename, std::ios::binary);
    Serialize(out);
}
void RuntimeNet::Deserialize(const std::string& filename)
{
    std::ifstream in(filename, std::ios::binary);
    Deserialize(in);
}

}  // namespace NetPimpl
