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

#include <cmath>
#include <deque>
#include <filesystem>
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

// One computational layer of the trainer's solver stack (dnn_trainer keeps one dlib::sgd per add_layer, output side first)
struct SolverSlot { int dlib_layer; size_t params; };

struct TrainingNet::Impl {
    ab_net* net = nullptr;
    bool allocated = false;          // parameters exist (first StartTraining, Deserialize, or a loaded synchronization file)
    bool step_pending = false;
    // configuration collected before the parameters are allocated
    int class_count = 2;
    double width_scaler = 1.0;
    int min_filter_count = 1;
    float weight_decay = 0.001f, momentum = 0.9f;
    unsigned long bn_window = 100;
    bool bn_window_dirty = false;    // SetAllBatchNormalizationRunningStatsWindowSizes was called since the windows were last pushed
    // ---- dnn_trainer's own state, member for member (trainer.h:1316-1372; defaults from init(), trainer.h:886-921)
    dlib::running_stats<double> rs;
    dlib::running_stats_decayed<double> rs_test = dlib::running_stats_decayed<double>(200);
    std::deque<double> previous_loss_values, test_previous_loss_values, previous_loss_values_to_keep_until_disk_sync;
    unsigned long max_num_epochs = 10000;
    size_t mini_batch_size = 128;
    bool verbose = false;
    double learning_rate = 1e-2, max_learning_rate_to_try_if_no_progress = 1e-2, min_learning_rate = 1e-5;
    unsigned long iter_without_progress_thresh = 2000, steps_without_progress = 0;
    unsigned long test_iter_without_progress_thresh = 500, test_steps_without_progress = 0;
    double learning_rate_shrink = 0.1;
    unsigned long epoch_iteration = 0;
    size_t epoch_pos = 0;
    unsigned long long train_one_step_calls = 0, test_one_step_calls = 0;
    dlib::matrix<double, 0, 1> lr_schedule;
    long lr_schedule_pos = 0;
    unsigned long gradient_check_budget = 0;
    size_t main_iteration_counter = 0, main_iteration_counter_at_last_disk_sync = 0;
    double prob_loss_increasing_thresh_default_value = 0.99, prob_loss_increasing_thresh_max_value = 0.99999, prob_loss_increasing_thresh = 0.99;
    bool updated_net_since_last_sync = false;
    unsigned long previous_loss_values_dump_amount = 400, test_previous_loss_values_dump_amount = 100;
    std::string sync_filename;
    std::chrono::seconds time_between_syncs = std::chrono::minutes(15);
    std::chrono::time_point<std::chrono::system_clock> last_sync_time = std::chrono::system_clock::now(), last_print = std::chrono::system_clock::now();
    // staging
    std::vector<uint8_t> images;
    std::vector<ab_wlabel> labels;

    ~Impl() { if (net) ab_net_destroy(net); }

    void push_hyper()
    {
        check(ab_set_hyper(net, AB_HYPER_LEARNING_RATE, learning_rate));
        check(ab_set_hyper(net, AB_HYPER_WEIGHT_DECAY, weight_decay));
        check(ab_set_hyper(net, AB_HYPER_MOMENTUM, momentum));
        // the windows live in the net (bn_::running_stats_window_size is serialized with it): a loaded net / state file keeps its
        // own until the setter is called again
        if (bn_window_dirty) { check(ab_set_hyper(net, AB_HYPER_BN_STATS_WINDOW, (double)bn_window)); bn_window_dirty = false; }
    }

    // the solver stack: every dlib layer that is an add_layer (not loss, tag, input), output side first (core.h:3806-3816)
    std::vector<SolverSlot> solver_slots() const
    {
        int n = 0;
        check(ab_get_layer_types(net, nullptr, 0, &n));
        std::vector<int32_t> kinds(n);
        check(ab_get_layer_types(net, kinds.data(), n, &n));
        std::vector<SolverSlot> slots;
        for (int i = 0; i < n; ++i) {
            if (kinds[i] == AB_LAYER_LOSS || kinds[i] == AB_LAYER_TAG || kinds[i] == AB_LAYER_INPUT) continue;
            size_t cnt = 0;
            if (kinds[i] == AB_LAYER_CON || kinds[i] == AB_LAYER_CONT || kinds[i] == AB_LAYER_BN) check(ab_get_layer_params(net, i, nullptr, 0, &cnt));
            slots.push_back(SolverSlot{i, cnt});
        }
        return slots;
    }

    void drop_some_previous_loss_values()   // trainer.h:1029-1033
    {
        for (unsigned long cnt = 0; cnt < previous_loss_values_dump_amount + iter_without_progress_thresh / 10 && !previous_loss_values.empty(); ++cnt)
            previous_loss_values.pop_front();
    }
    void drop_some_test_previous_loss_values()   // trainer.h:1036-1040
    {
        for (unsigned long cnt = 0; cnt < test_previous_loss_values_dump_amount + test_iter_without_progress_thresh / 10 && !test_previous_loss_values.empty(); ++cnt)
            test_previous_loss_values.pop_front();
    }

    // what dnn_trainer's thread does after it has taken a training job (trainer.h:743-864): record_loss, [update -- done on
    // the device by the step itself], learning-rate shrink on a plateau, learning-rate schedule
    void finish_step()
    {
        if (!step_pending) return;
        double loss = 0;
        check(ab_train_sync(net, &loss));
        step_pending = false;
        updated_net_since_last_sync = true;
        ++main_iteration_counter;
        gradient_check_budget += 200;                       // record_loss, trainer.h:612-628
        rs.add(loss);
        previous_loss_values.push_back(loss);
        while (previous_loss_values.size() > iter_without_progress_thresh) previous_loss_values.pop_front();
        if (!sync_filename.empty()) previous_loss_values_to_keep_until_disk_sync.push_back(loss);
        if (gradient_check_budget > iter_without_progress_thresh && learning_rate_shrink != 1) {   // trainer.h:835-864
            gradient_check_budget = 0;
            steps_without_progress = dlib::count_steps_without_decrease(previous_loss_values);
            if (steps_without_progress >= iter_without_progress_thresh) {
                steps_without_progress = dlib::count_steps_without_decrease_robust(previous_loss_values);
                if (steps_without_progress >= iter_without_progress_thresh) {
                    learning_rate = learning_rate_shrink * learning_rate;
                    steps_without_progress = 0;
                    drop_some_previous_loss_values();
                }
            }
        } else if (lr_schedule.size() != 0) {
            if (lr_schedule_pos < lr_schedule.size()) learning_rate = lr_schedule(lr_schedule_pos++);
            else learning_rate = lr_schedule(lr_schedule.size() - 1) * 0.99;
        }
        if (verbose) {   // print_periodic_verbose_status, trainer.h:1283-1310: every 20 s
            const auto now = std::chrono::system_clock::now();
            if (now - last_print > std::chrono::seconds(20)) {
                last_print = now;
                std::cout << "step#: " << train_one_step_calls << "  learning rate: " << learning_rate << "  average loss: " << rs.mean()
                          << "  steps without apparent progress: " << steps_without_progress << std::endl;
                rs.clear();
            }
        }
    }

    // ---- the synchronization file: dlib's dnn_trainer version 13 stream, field for field (trainer.h:931-964) ----------
    void write_state(std::ostream& out)
    {
        const int version = 13;
        dlib::serialize(version, out);
        int nl = 0;
        check(ab_get_layer_types(net, nullptr, 0, &nl));
        const size_t num_layers = (size_t)nl;            // net_type::num_layers (231 for LEVEL_COUNT = 6)
        dlib::serialize(num_layers, out);
        dlib::serialize(rs, out);
        dlib::serialize(rs_test, out);
        dlib::serialize(previous_loss_values, out);
        dlib::serialize(max_num_epochs, out);
        dlib::serialize(mini_batch_size, out);
        dlib::serialize(verbose, out);
        save_to_stream(net, 0, out);                       // serialize(item.net): the clean()ed net_type
        const std::vector<SolverSlot> slots = solver_slots();
        const unsigned long nsolvers = (unsigned long)slots.size();
        dlib::serialize(nsolvers, out);                    // std::vector<sgd>
        std::vector<float> v;
        for (const SolverSlot& s : slots) {
            dlib::serialize(std::string("sgd2"), out);     // solvers.h:121-127
            dlib::resizable_tensor t;
            if (s.params) {
                v.resize(s.params);
                size_t n = 0;
                check(ab_get_layer_momentum(net, s.dlib_layer, v.data(), v.size(), &n));
                t.set_size((long long)s.params);
                std::memcpy(t.host_write_only(), v.data(), s.params * sizeof(float));
            }
            dlib::serialize(t, out);
            dlib::serialize(weight_decay, out);
            dlib::serialize(momentum, out);
        }
        dlib::serialize(learning_rate, out);
        dlib::serialize(min_learning_rate, out);
        dlib::serialize(iter_without_progress_thresh, out);
        dlib::serialize(steps_without_progress, out);
        dlib::serialize(learning_rate_shrink, out);
        dlib::serialize(epoch_iteration, out);
        dlib::serialize(epoch_pos, out);
        dlib::serialize(train_one_step_calls, out);
        dlib::serialize(test_one_step_calls, out);
        dlib::serialize(lr_schedule, out);
        dlib::serialize(lr_schedule_pos, out);
        dlib::serialize(test_iter_without_progress_thresh, out);
        dlib::serialize(test_steps_without_progress, out);
        dlib::serialize(test_previous_loss_values, out);
        dlib::serialize(previous_loss_values_dump_amount, out);
        dlib::serialize(test_previous_loss_values_dump_amount, out);
        dlib::serialize(previous_loss_values_to_keep_until_disk_sync, out);
    }

    void read_state(std::istream& in)
    {
        int version = 0;
        dlib::deserialize(version, in);
        if (version != 13) throw dlib::serialization_error("Unexpected version found while deserializing dlib::dnn_trainer.");
        size_t num_layers = 0;
        dlib::deserialize(num_layers, in);
        dlib::deserialize(rs, in);
        dlib::deserialize(rs_test, in);
        dlib::deserialize(previous_loss_values, in);
        dlib::deserialize(max_num_epochs, in);
        dlib::deserialize(mini_batch_size, in);
        dlib::deserialize(verbose, in);
        load_from_stream(net, in);
        allocated = true;
        int nl = 0;
        check(ab_get_layer_types(net, nullptr, 0, &nl));
        if (num_layers != (size_t)nl) {
            std::ostringstream sout;
            sout << "Error deserializing dlib::dnn_trainer.  The saved sync file is for a network with a different number of layers.  We expected "
                 << nl << " but instead the file contains " << num_layers << " layers.";
            throw dlib::serialization_error(sout.str());
        }
        const std::vector<SolverSlot> slots = solver_slots();
        unsigned long nsolvers = 0;
        dlib::deserialize(nsolvers, in);
        if (nsolvers != slots.size()) throw dlib::serialization_error("dnn_trainer sync file: unexpected number of solvers");
        for (const SolverSlot& s : slots) {
            std::string tag;
            dlib::deserialize(tag, in);
            if (tag != "sgd2") throw dlib::serialization_error("Unexpected version found while deserializing dlib::sgd.");
            dlib::resizable_tensor t;
            dlib::deserialize(t, in);
            dlib::deserialize(weight_decay, in);
            dlib::deserialize(momentum, in);
            if (t.size() != 0) {   // (empty before the layer's first update: sgd allocates v lazily, solvers.h:44-49)
                if (t.size() != s.params) throw dlib::serialization_error("dnn_trainer sync file: solver state size does not match layer<" + std::to_string(s.dlib_layer) + ">");
                check(ab_set_layer_momentum(net, s.dlib_layer, t.host(), t.size()));
            }
        }
        unsigned long ltemp = 0;
        dlib::deserialize(learning_rate, in);
        dlib::deserialize(min_learning_rate, in);
        dlib::deserialize(ltemp, in); iter_without_progress_thresh = ltemp;
        dlib::deserialize(ltemp, in); steps_without_progress = ltemp;
        dlib::deserialize(learning_rate_shrink, in);
        dlib::deserialize(epoch_iteration, in);
        dlib::deserialize(epoch_pos, in);
        dlib::deserialize(train_one_step_calls, in);
        dlib::deserialize(test_one_step_calls, in);
        dlib::deserialize(lr_schedule, in);
        dlib::deserialize(lr_schedule_pos, in);
        dlib::deserialize(ltemp, in); test_iter_without_progress_thresh = ltemp;
        dlib::deserialize(ltemp, in); test_steps_without_progress = ltemp;
        dlib::deserialize(test_previous_loss_values, in);
        dlib::deserialize(previous_loss_values_dump_amount, in);
        dlib::deserialize(test_previous_loss_values_dump_amount, in);
        dlib::deserialize(previous_loss_values_to_keep_until_disk_sync, in);
        max_learning_rate_to_try_if_no_progress = learning_rate;
        push_hyper();
    }

    static bool exists(const std::string& f) { return (bool)std::ifstream(f, std::ios::binary); }
    static std::filesystem::file_time_type mtime(const std::string& f) { std::error_code ec; return std::filesystem::last_write_time(f, ec); }
    std::string newest_syncfile() const   // select_newest_file(sync_filename, sync_filename + "_"), trainer.h:1116-1120
    {
        const std::string a = sync_filename, b = sync_filename + "_";
        if (!exists(b)) return a;
        if (!exists(a)) return b;
        return mtime(a) >= mtime(b) ? a : b;
    }
    std::string oldest_syncfile() const   // select_oldest_file, trainer.h:1122-1126: a file that does not exist yet is the oldest
    {
        const std::string a = sync_filename, b = sync_filename + "_";
        if (!exists(a)) return a;
        if (!exists(b)) return b;
        return mtime(a) <= mtime(b) ? a : b;
    }

    bool loss_increased_since_last_disk_sync()   // trainer.h:1128-1175
    {
        const size_t gradient_updates_since_last_sync = main_iteration_counter - main_iteration_counter_at_last_disk_sync;
        if (!exists(newest_syncfile())) return false;
        while (previous_loss_values_to_keep_until_disk_sync.size() > 2 * gradient_updates_since_last_sync)
            previous_loss_values_to_keep_until_disk_sync.pop_front();
        for (double x : previous_loss_values_to_keep_until_disk_sync)
            if (std::isnan(x)) return true;
        if (gradient_updates_since_last_sync < 30) return false;
        const double prob1 = dlib::probability_values_are_increasing(previous_loss_values_to_keep_until_disk_sync);
        const double prob2 = dlib::probability_values_are_increasing_robust(previous_loss_values_to_keep_until_disk_sync);
        if (std::max(prob1, prob2) > prob_loss_increasing_thresh) {
            prob_loss_increasing_thresh = std::min(0.1 * prob_loss_increasing_thresh + 0.9 * 1, prob_loss_increasing_thresh_max_value);
            return true;
        }
        prob_loss_increasing_thresh = std::max(std::pow(prob_loss_increasing_thresh, 10.0), prob_loss_increasing_thresh_default_value);
        return false;
    }

    void sync_to_disk(bool do_it_now = false)   // trainer.h:1043-1113
    {
        if (!updated_net_since_last_sync && !step_pending) return;
        if (sync_filename.empty()) return;
        if (!(std::chrono::system_clock::now() - last_sync_time > time_between_syncs || do_it_now)) return;
        finish_step();   // wait_for_thread_to_pause
        if (!updated_net_since_last_sync) return;
        if (loss_increased_since_last_disk_sync()) {
            const double temp = max_learning_rate_to_try_if_no_progress;
            std::ifstream fin(newest_syncfile(), std::ios::binary);
            read_state(fin);
            if (verbose) std::cout << "Loss has been increasing, reloading saved state from " << newest_syncfile() << std::endl;
            if (prob_loss_increasing_thresh >= prob_loss_increasing_thresh_max_value) {
                if (verbose) std::cout << "(and while at it, also shrinking the learning rate)" << std::endl;
                max_learning_rate_to_try_if_no_progress = temp * learning_rate_shrink;
                learning_rate = std::min(learning_rate_shrink * learning_rate, max_learning_rate_to_try_if_no_progress);
                steps_without_progress = 0;
                test_steps_without_progress = 0;
                drop_some_previous_loss_values();
                drop_some_test_previous_loss_values();
            }
        } else {
            const std::string filename = oldest_syncfile();
            {
                std::ofstream fout(filename, std::ios::binary);
                if (!fout) throw dlib::serialization_error("Unable to open " + filename + " for writing.");
                write_state(fout);
            }
            if (verbose) std::cout << "Saved state to " << filename << std::endl;
        }
        last_sync_time = std::chrono::system_clock::now();
        main_iteration_counter_at_last_disk_sync = main_iteration_counter;
        updated_net_since_last_sync = false;
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
    if (pimpl->net) pimpl->finish_step();   // get_net(force_flush_to_disk::no): may block (NetPimpl.cpp:30-32)
    delete pimpl;
    pimpl = new Impl();
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
    // dnn_trainer::set_learning_rate (trainer.h:455-472): waits for the thread, drops the schedule, resets the progress counters
    pimpl->finish_step();
    if (pimpl->learning_rate != learningRate) {
        pimpl->steps_without_progress = 0;
        pimpl->test_steps_without_progress = 0;
        pimpl->previous_loss_values.clear();
        pimpl->test_previous_loss_values.clear();
        pimpl->previous_loss_values_to_keep_until_disk_sync.clear();
    }
    pimpl->learning_rate = learningRate;
    pimpl->max_learning_rate_to_try_if_no_progress = learningRate;
    pimpl->lr_schedule.set_size(0);
}
void TrainingNet::SetIterationsWithoutProgressThreshold(unsigned long threshold)
{
    pimpl->finish_step();
    pimpl->lr_schedule.set_size(0);
    pimpl->iter_without_progress_thresh = threshold;
}
void TrainingNet::SetPreviousLossValuesDumpAmount(unsigned long dump_amount) { pimpl->previous_loss_values_dump_amount = dump_amount; }
void TrainingNet::SetAllBatchNormalizationRunningStatsWindowSizes(unsigned long window_size)
{
    if (window_size == 0) throw dlib::fatal_error(dlib::EBROKEN_ASSERT, "The batch normalization running stats window size can't be 0.");
    pimpl->bn_window = window_size;
    pimpl->bn_window_dirty = true;
    if (pimpl->allocated) pimpl->push_hyper();
}
void TrainingNet::SetLearningRateShrinkFactor(double f)
{
    pimpl->finish_step();
    pimpl->lr_schedule.set_size(0);
    pimpl->learning_rate_shrink = f;
    pimpl->steps_without_progress = 0;
    pimpl->test_steps_without_progress = 0;
}
// dnn_trainer::set_synchronization_file (trainer.h:411-424): remember the name and load the newest existing file -- the same
// version-13 stream the reference writes, so a run can be resumed across the two implementations (annonet_train_main.cpp:403)
void TrainingNet::SetSynchronizationFile(const std::string& filename, std::chrono::seconds time_between_syncs)
{
    if (!pimpl->net) Initialize();
    pimpl->last_sync_time = std::chrono::system_clock::now();
    pimpl->sync_filename = filename;
    pimpl->time_between_syncs = time_between_syncs;
    std::ifstream fin(pimpl->newest_syncfile(), std::ios::binary);
    if (fin) {
        pimpl->finish_step();
        pimpl->read_state(fin);
    }
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
    // Stage this mini-batch on the host WHILE the previous step still runs on the device: dnn_trainer converts step t + 1's
    // input (to_tensor) on the caller's thread before it hands the job to the trainer thread (trainer.h:216-224,1206-1253).
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
        std::memcpy(pimpl->labels.data() + i * px, &training_labels[i](0, 0), px * sizeof(ab_wlabel));   // dlib matrices are row-major and dense
    }
    pimpl->sync_to_disk();   // train_one_step (trainer.h:216-224): sync_to_disk(); send_job(); ++train_one_step_calls
    pimpl->finish_step();    // rendezvous: the trainer takes the next job when the previous step is done (trainer.h:1252)
    if (!pimpl->allocated) {
        check(ab_net_init_random(pimpl->net, pimpl->class_count, pimpl->width_scaler, pimpl->min_filter_count, (uint64_t)std::rand()));
        pimpl->allocated = true;
        pimpl->bn_window_dirty = true;   // a fresh net: bn_ default window is 100 (layers.h:1298), the trainer's setter value applies
    }
    pimpl->push_hyper();
    check(ab_train_step_async(pimpl->net, pimpl->images.data(), pimpl->labels.data(), (int)n, (int)h, (int)w));
    pimpl->step_pending = true;
    ++pimpl->train_one_step_calls;
}

// dnn_trainer::get_learning_rate reads an atomic the trainer thread writes after each step -- it does NOT wait for the running
// step (trainer.h:474-478).  Same here: the value the last finished step left behind.
double TrainingNet::GetLearningRate() const { return pimpl->learning_rate; }

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
    t->finish_step();        // get_net(): wait_for_thread_to_pause ...
    t->sync_to_disk(true);   // ... and sync_to_disk(force_flush_to_disk::yes) (trainer.h:143-151, NetPimpl.cpp:226-232)
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
