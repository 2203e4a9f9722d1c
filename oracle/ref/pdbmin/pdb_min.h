/* pdb_min.h -- TEST INFRASTRUCTURE (part of the oracle).  A minimal stand-in for the slice of netsDB's object model and
 * lambda API that the reference's UDF headers use, so that those headers -- src/FF/headers/*.h, src/LSTM/headers/*.h,
 * src/conv2d_memory_fusion/headers/*.h, src/sharedLibraries/headers/{MatrixBlock,LASilly*Multiply*}.h -- compile
 * UNMODIFIED, from where they lie under /root/reference, into oracle/_ref/libff_ref.so (oracle/Makefile, target `ref`).
 * The real object model cannot be built here (it needs SCons-generated sources).
 *
 * What is stood in: Object, Handle (shared ownership instead of allocation blocks), makeObject, Vector, Map, String,
 * Hasher, the Computation base classes, and Lambdas that are EVALUATED on the handles they were created from (the
 * reference compiles them into pipelines; here getSelection / getProjection are simply called per tuple).
 * What is NOT stood in: any arithmetic.  Every number libff_ref.so produces comes out of the reference's own lambda
 * bodies and operator+ implementations.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <type_traits>
#include <vector>

using namespace std; /* the reference's headers rely on it (unqualified cout / endl in FFOutputLayer.h:38-39) */

namespace pdb {

class Object {
public:
    virtual ~Object() {}
    virtual size_t hash() const { return 0; }
};
typedef Object PDBObject;

struct NotEnoughSpace : public std::exception {};

#define ENABLE_DEEP_COPY
#ifndef GET_V_TABLE
#define GET_V_TABLE(T)
#endif

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
    /* Handle.h: two handles are equal when their targets are (operator== of the target type) */
    bool operator==(const Handle<T> &o) const { return p && o.p ? (*p == *o.p) : (!p && !o.p); }
    Handle &operator=(std::nullptr_t) { p.reset(); return *this; }
    const std::shared_ptr<T> &ptr() const { return p; }
    bool isNullPtr() const { return !p; }

private:
    std::shared_ptr<T> p;
};
template <class T, class... A> Handle<T> makeObject(A &&...a) { return Handle<T>(std::make_shared<T>(std::forward<A>(a)...)); }
template <class To, class From> Handle<To> unsafeCast(const Handle<From> &h) { return Handle<To>(std::dynamic_pointer_cast<To>(h.ptr())); }

template <class T> struct Hasher {
    static size_t hash(const T &x) { return std::hash<T>()(x); }
};

/* PDBVector.h / Array.cc:84-100: (capacity) or (capacity, used) constructors, the latter zero-fills */
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

/* PDBMap.h: operator[], count, iteration over entries with .key / .value */
template <class K, class V>
class Map : public Object {
public:
    struct Entry {
        K key;
        V value;
    };
    V &operator[](const K &k) {
        for (auto &e : items) if (e->key == k) return e->value;
        items.push_back(std::unique_ptr<Entry>(new Entry{k, V()}));
        return items.back()->value;
    }
    int count(const K &k) {
        for (auto &e : items) if (e->key == k) return 1;
        return 0;
    }
    size_t size() const { return items.size(); }
    struct Iter {
        typename std::vector<std::shared_ptr<Entry>>::iterator it;
        bool operator!=(const Iter &o) const { return it != o.it; }
        void operator++() { ++it; }
        Entry &operator*() { return **it; }
    };
    Iter begin() { return Iter{items.begin()}; }
    Iter end() { return Iter{items.end()}; }

private:
    std::vector<std::shared_ptr<Entry>> items;
};

typedef std::string String;
struct UseTemporaryAllocationBlock {
    explicit UseTemporaryAllocationBlock(size_t) {}
};

/* ---- lambdas, evaluated in place ------------------------------------------------------------------ */
template <class Out>
class Lambda {
public:
    Lambda() {}
    explicit Lambda(std::function<Out()> f) : fn(std::move(f)) {}
    Out eval() const { return fn(); }

private:
    std::function<Out()> fn;
};
/* the reference's lambda bodies take their inputs as non-const Handle references */
template <class In1, class F> auto makeLambda(Handle<In1> a, F f) -> Lambda<decltype(f(a))> {
    return Lambda<decltype(f(a))>([a, f]() mutable { return f(a); });
}
template <class In1, class In2, class F> auto makeLambda(Handle<In1> a, Handle<In2> b, F f) -> Lambda<decltype(f(a, b))> {
    return Lambda<decltype(f(a, b))>([a, b, f]() mutable { return f(a, b); });
}
template <class In1, class In2, class In3, class F>
auto makeLambda(Handle<In1> a, Handle<In2> b, Handle<In3> c, F f) -> Lambda<decltype(f(a, b, c))> {
    return Lambda<decltype(f(a, b, c))>([a, b, c, f]() mutable { return f(a, b, c); });
}
#define makeLambdaFromMethod(VAR, METHOD)                                                      \
    (::pdb::Lambda<typename std::decay<decltype((VAR)->METHOD())>::type>(                      \
        [VAR]() -> typename std::decay<decltype((VAR)->METHOD())>::type { return (VAR)->METHOD(); }))
#define makeLambdaFromMember(VAR, MEMBER)                                                      \
    (::pdb::Lambda<typename std::decay<decltype((VAR)->MEMBER)>::type>(                        \
        [VAR]() -> typename std::decay<decltype((VAR)->MEMBER)>::type { return (VAR)->MEMBER; }))
template <class A, class B> Lambda<bool> operator==(const Lambda<A> &l, const Lambda<B> &r) {
    return Lambda<bool>([l, r]() { return l.eval() == r.eval(); });
}
inline Lambda<bool> operator&&(const Lambda<bool> &l, const Lambda<bool> &r) {
    return Lambda<bool>([l, r]() { return l.eval() && r.eval(); });
}

/* ---- computation base classes (only what the UDF headers override) ----------------------------------- */
class Computation : public Object {
public:
    virtual std::string getComputationType() { return "Computation"; }
    bool setInput(Handle<Computation>) { return true; }
    bool setInput(int, Handle<Computation>) { return true; }
};
template <class Out, class In1, class In2, class In3 = void>
class JoinComp : public Computation {
public:
    virtual Lambda<bool> getSelection(Handle<In1> in1, Handle<In2> in2, Handle<In3> in3) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In1> in1, Handle<In2> in2, Handle<In3> in3) = 0;
};
template <class Out, class In1, class In2>
class JoinComp<Out, In1, In2, void> : public Computation {
public:
    virtual Lambda<bool> getSelection(Handle<In1> in1, Handle<In2> in2) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In1> in1, Handle<In2> in2) = 0;
};
template <class Out, class In, class Key, class Val>
class ClusterAggregateComp : public Computation {
public:
    virtual Lambda<Key> getKeyProjection(Handle<In> aggMe) = 0;
    virtual Lambda<Val> getValueProjection(Handle<In> aggMe) = 0;
};
template <class Out, class In>
class SelectionComp : public Computation {
public:
    virtual Lambda<bool> getSelection(Handle<In> in) = 0;
    virtual Lambda<Handle<Out>> getProjection(Handle<In> in) = 0;
};
template <class Out, class In>
class MultiSelectionComp : public Computation {
public:
    virtual Lambda<bool> getSelection(Handle<In> in) = 0;
    virtual Lambda<Vector<Handle<Out>>> getProjection(Handle<In> in) = 0;
};
template <class T>
class ScanUserSet : public Computation {
public:
    void setDatabaseName(std::string) {}
    void setSetName(std::string) {}
};
template <class T>
class WriteUserSet : public Computation {
public:
    void setOutput(std::string, std::string) {}
};
class ExportableObject : public Object {
public:
    virtual std::string toString() { return ""; }
    virtual void print() {}
    virtual std::string toSchemaString(std::string) { return ""; }
    virtual std::string toValueString(std::string) { return ""; }
    virtual std::vector<std::string> getSupportedFormats() { return {}; }
};
class StringIntPair : public Object {};

}  // namespace pdb
