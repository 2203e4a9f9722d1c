// pdb_shim.h -- the slice of netsDB's client / Computation API that the FF, conv2d and word2vec
// drivers touch, re-implemented over the C ABI of libnetsdb_b200.so so that
// driver code written against netsDB (src/tests/source/FFTest.cc, src/FF/source/SimpleFF.cc style)
// compiles and runs unchanged against the device.
//
// What is mirrored (same names, argument meaning, error behaviour):
//   object model   pdb::Object, Handle<T>, makeObject<T>, Vector<T>, String, makeObjectAllocatorBlock,
//                  UseTemporaryAllocationBlock, NotEnoughSpace, ENABLE_DEEP_COPY, GET_V_TABLE
//                  (objectModel/headers/Handle.h, builtInPDBObjects/headers/PDBVector.h)
//   lambdas        Lambda<T>, makeLambda, makeLambdaFromMethod, ==, &&   (lambdas/headers/LambdaCreationFunctions.h)
//   computations   Computation::setInput, JoinComp, ClusterAggregateComp, SelectionComp, MultiSelectionComp,
//                  ScanUserSet, WriteUserSet   (lambdas/headers/JoinComp.h:44-45,197-204;
//                  queryExecution/headers/ClusterAggregateComp.h:40-63; builtInPDBObjects/headers/ScanUserSet.h,
//                  WriteUserSet.h)
//   client         PDBClient, CatalogClient, PDBLogger   (mainClient/headers/PDBClient.h:40-259)
//
// There is no generic relational executor here and no CPU fallback: executeComputations lowers the
// graphs of this path to calls into the CUDA library (ff_stage.h) and rejects anything else.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <typeinfo>
#include <utility>
#include <vector>

#include "netsdb_b200.h"

namespace pdb {

// ---- object model ---------------------------------------------------------------------------------
class Object {
public:
    virtual ~Object() {}
    virtual size_t hash() const { return 0; }
};

struct NotEnoughSpace : public std::exception {
    const char *what() const noexcept override { return "NotEnoughSpace"; }
};

#define ENABLE_DEEP_COPY
#define GET_V_TABLE(T)

template <class T>
class Handle {
public:
    Handle() {}
    Handle(std::nullptr_t) {}
    explicit Handle(std::shared_ptr<T> q) : p(std::move(q)) {}
    template <class U, class = typename std::enable_if<std::is_convertible<U *, T *>::value>::type>
    Handle(const Handle<U> &o) : p(o.ptr()) {}
    T *operator->() const { return p.get(); }
    T &operator*() const { return *p; }
    bool operator==(std::nullptr_t) const { return !p; }
    bool operator!=(std::nullptr_t) const { return !!p; }
    template <class U> bool operator==(const Handle<U> &o) const { return p.get() == o.ptr().get(); }
    Handle &operator=(std::nullptr_t) { p.reset(); return *this; }
    const std::shared_ptr<T> &ptr() const { return p; }
    bool isNullPtr() const { return !p; }
    // distance from the handle to its target (objectModel/headers/Handle.h:211); only ever printed by the drivers
    int64_t getOffset() const { return reinterpret_cast<const char *>(p.get()) - reinterpret_cast<const char *>(this); }

private:
    std::shared_ptr<T> p;
};
template <class To, class From> Handle<To> unsafeCast(const Handle<From> &h) {
    return Handle<To>(std::dynamic_pointer_cast<To>(h.ptr()));
}
template <class T, class... A> Handle<T> makeObject(A &&...a) {
    return Handle<T>(std::make_shared<T>(std::forward<A>(a)...));
}

// pdb::Vector (builtInPDBObjects/headers/PDBVector.h): (capacity) or (capacity, used) constructors, the
// latter zero-fills `used` slots (Array.cc:84-100); resize() changes capacity only.
template <class T>
class Vector : public Object {
public:
    Vector() {}
    explicit Vector(uint32_t cap) { v.reserve(cap); }
    Vector(uint32_t cap, uint32_t used) { v.reserve(cap); v.assign(used, T()); }
    size_t size() const { return v.size(); }
    T &operator[](uint32_t i) { return v[i]; }
    const T &operator[](uint32_t i) const { return v[i]; }
    void push_back(const T &x) { v.push_back(x); }
    void resize(uint32_t cap) { v.reserve(cap); }
    void clear() { v.clear(); }
    void fill(const T &x) { std::fill(v.begin(), v.end(), x); }
    T *c_ptr() { return v.data(); }

private:
    std::vector<T> v;
};

// pdb::Map (builtInPDBObjects/headers/PDBMap.h): the slice the conv UDF types use
template <class K, class V>
class Map : public Object {
public:
    V &operator[](const K &k) { return m[k]; }
    int count(const K &k) { return static_cast<int>(m.count(k)); }
    size_t size() const { return m.size(); }
    typename std::map<K, V>::iterator begin() { return m.begin(); }
    typename std::map<K, V>::iterator end() { return m.end(); }

private:
    std::map<K, V> m;
};

using String = std::string;
inline void makeObjectAllocatorBlock(size_t, bool) {}
struct UseTemporaryAllocationBlock {
    explicit UseTemporaryAllocationBlock(size_t) {}
};
class PDBLogger {
public:
    explicit PDBLogger(std::string n) : name(std::move(n)) {}
    std::string name;
};
typedef std::shared_ptr<PDBLogger> PDBLoggerPtr;

// ---- lambdas --------------------------------------------------------------------------------------
// The planner only needs to know which key methods a selection compares; the projection callables are
// kept alive but never run per tuple: the join + aggregation pair is lowered as a whole.
template <class Out>
class Lambda {
public:
    Lambda() {}
    explicit Lambda(std::string d) : desc(std::move(d)) {}
    std::string describe() const { return desc; }
    std::shared_ptr<void> keep;
    std::string desc;
};
template <class In1, class F> auto makeLambda(Handle<In1> a, F f) -> Lambda<decltype(f(a))> {
    Lambda<decltype(f(a))> l("native_lambda");
    l.keep = std::make_shared<F>(std::move(f));
    return l;
}
template <class In1, class In2, class F> auto makeLambda(Handle<In1> a, Handle<In2> b, F f) -> Lambda<decltype(f(a, b))> {
    Lambda<decltype(f(a, b))> l("native_lambda");
    l.keep = std::make_shared<F>(std::move(f));
    return l;
}
template <class In1, class In2, class In3, class F>
auto makeLambda(Handle<In1> a, Handle<In2> b, Handle<In3> c, F f) -> Lambda<decltype(f(a, b, c))> {
    Lambda<decltype(f(a, b, c))> l("native_lambda");
    l.keep = std::make_shared<F>(std::move(f));
    return l;
}
#define makeLambdaFromMethod(VAR, METHOD) \
    (::pdb::Lambda<typename std::decay<decltype((VAR)->METHOD())>::type>(std::string("methodCall(") + #METHOD + ")"))
#define makeLambdaFromMember(VAR, MEMBER) \
    (::pdb::Lambda<typename std::decay<decltype((VAR)->MEMBER)>::type>(std::string("attAccess(") + #MEMBER + ")"))
template <class A, class B> Lambda<bool> operator==(const Lambda<A> &l, const Lambda<B> &r) {
    return Lambda<bool>("(" + l.desc + " == " + r.desc + ")");
}
inline Lambda<bool> operator&&(const Lambda<bool> &l, const Lambda<bool> &r) {
    return Lambda<bool>("(" + l.desc + " && " + r.desc + ")");
}
template <class A> Lambda<bool> operator&&(const Lambda<bool> &l, const Lambda<A> &r) {
    return Lambda<bool>("(" + l.desc + " && " + r.desc + ")");
}

// ---- computations ---------------------------------------------------------------------------------
class Computation : public Object {
public:
    virtual std::string getComputationType() { return "Computation"; }
    bool setInput(Handle<Computation> in) { return setInput(0, in); }
    bool setInput(int slot, Handle<Computation> in) {
        if (static_cast<int>(inputs.size()) <= slot) inputs.resize(slot + 1);
        inputs[slot] = in;
        if (in != nullptr) in->setNumConsumers(in->getNumConsumers() + 1);  // lambdas/headers/Computation.h:182
        return true;
    }
    int getNumConsumers() { return numConsumers; }
    void setNumConsumers(int n) { numConsumers = n; }
    int getNumInputs() const { return static_cast<int>(inputs.size()); }
    Handle<Computation> getIthInput(int i) const { return i < static_cast<int>(inputs.size()) ? inputs[i] : Handle<Computation>(); }
    void setAllocatorPolicy(int) {}

private:
    std::vector<Handle<Computation>> inputs;
    int numConsumers = 0;
};

// JoinComp<Out, In1, In2[, In3]> (lambdas/headers/JoinComp.h:44-45: up to five inputs there; the drivers of this
// path use two and, for LSTMThreeWaySum, three)
template <class Out, class In1, class In2, class In3 = void>
class JoinComp : public Computation {
public:
    std::string getComputationType() override { return "JoinComp"; }
    virtual Lambda<bool> getSelection(Handle<In1> in1, Handle<In2> in2, Handle<In3> in3) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In1> in1, Handle<In2> in2, Handle<In3> in3) = 0;
};
template <class Out, class In1, class In2>
class JoinComp<Out, In1, In2, void> : public Computation {
public:
    std::string getComputationType() override { return "JoinComp"; }
    virtual Lambda<bool> getSelection(Handle<In1> in1, Handle<In2> in2) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In1> in1, Handle<In2> in2) = 0;
};
template <class Out, class In, class Key, class Val>
class ClusterAggregateComp : public Computation {
public:
    std::string getComputationType() override { return "ClusterAggregationComp"; }
    virtual Lambda<Key> getKeyProjection(Handle<In> aggMe) = 0;
    virtual Lambda<Val> getValueProjection(Handle<In> aggMe) = 0;
};
template <class Out, class In>
class SelectionComp : public Computation {
public:
    std::string getComputationType() override { return "SelectionComp"; }
    virtual Lambda<bool> getSelection(Handle<In> in) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In> in) = 0;
};
template <class Out, class In>
class MultiSelectionComp : public Computation {
public:
    std::string getComputationType() override { return "MultiSelectionComp"; }
    virtual Lambda<bool> getSelection(Handle<In> in) = 0;
    virtual Lambda<Vector<Handle<Out>>> getProjection(Handle<In> in) = 0;
};
class ScanSetBase : public Computation {
public:
    std::string getComputationType() override { return "ScanUserSet"; }
    void setDatabaseName(std::string d) { dbName = std::move(d); }
    void setSetName(std::string s) { setName = std::move(s); }
    void setBatchSize(int) {}
    std::string getDatabaseName() const { return dbName; }
    std::string getSetName() const { return setName; }

protected:
    std::string dbName, setName;
};
template <class T>
class ScanUserSet : public ScanSetBase {
public:
    ScanUserSet() {}
    ScanUserSet(std::string d, std::string s) { dbName = std::move(d); setName = std::move(s); }
};
class WriteSetBase : public Computation {
public:
    std::string getComputationType() override { return "WriteUserSet"; }
    void setOutput(std::string d, std::string s) { dbName = std::move(d); setName = std::move(s); }
    std::string getDatabaseName() const { return dbName; }
    std::string getSetName() const { return setName; }

protected:
    std::string dbName, setName;
};
template <class T>
class WriteUserSet : public WriteSetBase {
public:
    WriteUserSet() {}
    WriteUserSet(std::string d, std::string s) { setOutput(std::move(d), std::move(s)); }
};

}  // namespace pdb
