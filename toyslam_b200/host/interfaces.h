// interfaces.h -- the reference's seams for this path, restated so the shims below drop into its factories.
//   IGraph      remote/graph/IGraph.h:3-7          (empty polymorphic base)
//   ISolver<T>  remote/solver/ISolver.h:10         (Solve(H, b, dims); accepted and ignored on PIPELINE=gpu,
//                                                   as the reference already overrides it, main.cpp:21-25)
//   IOptimizer<T> remote/optimizer/IOptimizer.h:10-26  (iterations + solver, virtual Optimize(IGraph*))
#pragma once
#include <memory>

class IGraph { public: virtual ~IGraph() = default; };

template <typename T> class IMatrix;
template <typename T> class ISolver {
public:
    virtual ~ISolver() = default;
    virtual void Solve(IMatrix<T>* H, IMatrix<T>* b, unsigned dims) = 0;
};

template <typename T> class IOptimizer {
public:
    IOptimizer(const unsigned iterations, std::unique_ptr<ISolver<T>> solver) : iterations(iterations), solver(std::move(solver)) {}
    virtual ~IOptimizer() = default;
    virtual void Optimize(IGraph* graph) = 0;
protected:
    unsigned iterations;
    std::unique_ptr<ISolver<T>> solver;
};
