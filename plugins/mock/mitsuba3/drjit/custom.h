// MOCK of Dr.Jit's custom differentiable operation interface (upstream include/drjit/custom.h,
// Dr.Jit 0.4.x naming: dr::CustomOp<DiffType, Output, Input...>, dr::custom<Op>(inputs...)).
// Not Dr.Jit: there is no AD graph.  dr::custom<Op>() evaluates the operation eagerly and keeps
// the operation object on a "tape" (mock_tape()); a test plays the role of dr::backward() by
// setting the output gradient on the recorded operation and calling its backward().  What this
// lets the repository check is the plugin's own CustomOp subclass (plugins/mitsuba3/merl_eval_op.h):
// that it compiles against the interface and that its eval()/backward() marshal the right buffers
// into libmerl_b200.
#pragma once
#include <drjit/array.h>

#include <memory>
#include <tuple>
#include <vector>

namespace drjit {

struct CustomOpBase {
    virtual ~CustomOpBase() = default;
    virtual void forward() { throw std::runtime_error("drjit mock: CustomOp::forward() not implemented"); }
    virtual void backward() { throw std::runtime_error("drjit mock: CustomOp::backward() not implemented"); }
    virtual const char* name() const = 0;
};

inline std::vector<std::shared_ptr<CustomOpBase>>& mock_tape() {
    static std::vector<std::shared_ptr<CustomOpBase>> tape;
    return tape;
}

template <typename DiffType_, typename Output_, typename... Input> struct CustomOp : CustomOpBase {
    using DiffType = DiffType_;
    using Output = Output_;
    using Inputs = std::tuple<Input...>;

    virtual Output eval(const Input&... input) = 0;

    // ---- what a subclass uses inside forward() / backward() (upstream names) ----------------
    template <size_t I> const auto& value_in() const { return std::get<I>(m_value_in); }
    const Output& value_out() const { return m_value_out; }
    const Output& grad_out() const { return m_grad_out; }
    template <size_t I, typename T> void set_grad_in(const T& g) { std::get<I>(m_grad_in) = g; }

    // ---- mock only: the AD driver's side -----------------------------------------------------
    void mock_set_grad_out(const Output& g) { m_grad_out = g; }
    template <size_t I> const auto& mock_grad_in() const { return std::get<I>(m_grad_in); }
    void mock_record(const Input&... input) {
        m_value_in = Inputs(input...);
        m_value_out = eval(input...);
    }

private:
    Inputs m_value_in, m_grad_in;
    Output m_value_out, m_grad_out;
};

// upstream: evaluates Op::eval() on detached inputs and, when an input carries gradients, inserts
// the operation into the AD graph.  Mock: always records.
template <typename Op, typename... Args> typename Op::Output custom(const Args&... args) {
    auto op = std::make_shared<Op>();
    op->mock_record(args...);
    mock_tape().push_back(op);
    return op->value_out();
}

}  // namespace drjit
