// Package klm binds libklm.so (include/klm.h) for the k8s-llm-monitor Go host.
//
// NOT COMPILED IN THIS REPOSITORY'S BUILD ENVIRONMENT (no Go toolchain there);
// the same C ABI is exercised by the ctypes tests. See INTEGRATION.md for the
// call-site changes in internal/scheduler/controller.go and
// internal/metrics/manager.go that use it.
package klm

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -lklm -lcudart
#include "klm.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

// Node flags: which of buildCandidates' per-item filters a node passes
// (internal/scheduler/controller.go:190, :198).
const (
	NodeHasName  = 1
	NodeStatusOK = 2
)

// Handle drives one GPU. All methods are safe for concurrent use.
type Handle struct{ h *C.klm_handle }

func lastErr(what string) error {
	return fmt.Errorf("%s: %s", what, C.GoString(C.klm_last_error()))
}

// New opens device `device`. There is no CPU fallback: on a machine without a
// B200 this returns an error and the caller keeps its Go code path.
func New(device int) (*Handle, error) {
	dev := C.int(device)
	var h *C.klm_handle
	if rc := C.klm_create(&dev, 1, &h); rc != 0 {
		return nil, lastErr("klm_create")
	}
	return &Handle{h}, nil
}

func (k *Handle) Close() { C.klm_destroy(k.h); k.h = nil }

func f64p(s []float64) *C.double {
	if len(s) == 0 {
		return nil
	}
	return (*C.double)(unsafe.Pointer(&s[0]))
}
func f32p(s []float32) *C.float {
	if len(s) == 0 {
		return nil
	}
	return (*C.float)(unsafe.Pointer(&s[0]))
}
func i64p(s []int64) *C.int64_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.int64_t)(unsafe.Pointer(&s[0]))
}
func i32p(s []int32) *C.int32_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.int32_t)(unsafe.Pointer(&s[0]))
}
func u8p(s []uint8) *C.uint8_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint8_t)(unsafe.Pointer(&s[0]))
}

// NodesSet installs the UAV snapshot of one reconcile (controller.go:96-98).
func (k *Handle) NodesSet(battery []float64, flags []uint8) error {
	if len(battery) != len(flags) {
		return fmt.Errorf("klm: battery and flags differ in length")
	}
	if rc := C.klm_nodes_set(k.h, C.int64_t(len(battery)), f64p(battery), u8p(flags)); rc != 0 {
		return lastErr("klm_nodes_set")
	}
	return nil
}

// ScheduleBatch replaces buildCandidates + sort.SliceStable + candidates[0]
// (controller.go:152-161, :174-221) for a batch of independent requests.
// idx[r] == -1 means "no candidate" (controller.go:153-158).
func (k *Handle) ScheduleBatch(minBattery []float64, prefOff, prefIDs []int32) ([]int32, []float64, error) {
	r := len(minBattery)
	idx, score := make([]int32, r), make([]float64, r)
	if r == 0 {
		return idx, score, nil
	}
	if len(prefOff) != r+1 {
		return nil, nil, fmt.Errorf("klm: prefOff must have len(minBattery)+1 entries")
	}
	ids := prefIDs
	if len(ids) == 0 {
		ids = []int32{0}
	}
	rc := C.klm_schedule_batch(k.h, C.int64_t(r), f64p(minBattery), i32p(prefOff), i32p(ids),
		i32p(idx), f64p(score), C.KLM_ALGO_DECOMPOSED)
	if rc != 0 {
		return nil, nil, lastErr("klm_schedule_batch")
	}
	return idx, score, nil
}

// Reconcile is NodesSet + ScheduleBatch in one call (klm_reconcile): the whole of
// controller.go:88-110 for one cycle, one lock and -- for a small reconcile -- one upload, one
// download and one synchronisation.
func (k *Handle) Reconcile(battery []float64, flags []uint8, minBattery []float64, prefOff, prefIDs []int32) ([]int32, []float64, error) {
	r := len(minBattery)
	idx, score := make([]int32, r), make([]float64, r)
	rc := C.klm_reconcile(k.h, C.int64_t(len(battery)), f64p(battery), u8p(flags), C.int64_t(r),
		f64p(minBattery), i32p(prefOff), i32p(prefIDs), i32p(idx), f64p(score), C.KLM_ALGO_DECOMPOSED)
	if rc != 0 {
		return nil, nil, lastErr("klm_reconcile")
	}
	return idx, score, nil
}

// NodeCols are the columns calculateClusterMetrics reads (manager.go:499-535).
type NodeCols struct {
	CPUCapacity, CPUUsage, MemoryCapacity, MemoryUsage []int64
	GPUCount                                           []int32
	Healthy                                            []uint8
	GPUUsage                                           []float64
}

// ClusterOut is the numeric part of pkg/metrics.ClusterMetrics.
type ClusterOut struct {
	TotalNodes, HealthyNodes, TotalPods, RunningPods int64
	TotalCPU, UsedCPU, TotalMemory, UsedMemory       int64
	TotalGPUs, AvailableGPUs                         int64
	CPUUsageRate, MemoryUsageRate                    float64
	HealthStatus                                     int32  // 0 healthy, 1 warning, 2 critical
	Issues                                           uint32 // bit0 unhealthy nodes, bit1 high CPU, bit2 high memory
}

// ClusterAggregate replaces Manager.calculateClusterMetrics (manager.go:493-565).
func (k *Handle) ClusterAggregate(c *NodeCols, podRunning []uint8) (*ClusterOut, error) {
	var pin runtime.Pinner
	defer pin.Unpin()
	cc := c.toC(&pin)
	var o C.klm_cluster_out
	rc := C.klm_cluster_aggregate(k.h, &cc, C.int64_t(len(c.CPUCapacity)), u8p(podRunning),
		C.int64_t(len(podRunning)), &o)
	if rc != 0 {
		return nil, lastErr("klm_cluster_aggregate")
	}
	out := clusterFromC(&o)
	return &out, nil
}

// pinSlice pins the backing array of s for the duration of one call. A C struct that holds
// Go pointers may be passed to C only if those pointers are pinned (cgo rules, Go >= 1.21).
func pinSlice[T any](p *runtime.Pinner, s []T) {
	if len(s) > 0 {
		p.Pin(&s[0])
	}
}

func (c *NodeCols) toC(p *runtime.Pinner) C.klm_node_cols {
	pinSlice(p, c.CPUCapacity)
	pinSlice(p, c.CPUUsage)
	pinSlice(p, c.MemoryCapacity)
	pinSlice(p, c.MemoryUsage)
	pinSlice(p, c.GPUCount)
	pinSlice(p, c.Healthy)
	pinSlice(p, c.GPUUsage)
	return C.klm_node_cols{
		cpu_capacity: i64p(c.CPUCapacity), cpu_usage: i64p(c.CPUUsage),
		memory_capacity: i64p(c.MemoryCapacity), memory_usage: i64p(c.MemoryUsage),
		gpu_count: i32p(c.GPUCount), healthy: u8p(c.Healthy),
		gpu_usage: f64p(c.GPUUsage), n_gpu_usage: C.int64_t(len(c.GPUUsage)),
	}
}

func clusterFromC(o *C.klm_cluster_out) ClusterOut {
	return ClusterOut{
		int64(o.total_nodes), int64(o.healthy_nodes), int64(o.total_pods), int64(o.running_pods),
		int64(o.total_cpu), int64(o.used_cpu), int64(o.total_memory), int64(o.used_memory),
		int64(o.total_gpus), int64(o.available_gpus), float64(o.cpu_usage_rate),
		float64(o.memory_usage_rate), int32(o.health_status), uint32(o.issues)}
}

// FleetScan replaces GetHealthyUAVCount and GetLowBatteryUAVs
// (internal/metrics/sources/uav_metrics.go:197-233); indices come back ascending.
func (k *Handle) FleetScan(statusOK []uint8, remainingPercent []float64, threshold float64) (int64, []int32, error) {
	n := len(statusOK)
	low := make([]int32, n+1)
	var healthy, nLow C.int64_t
	rc := C.klm_fleet_scan(k.h, C.int64_t(n), u8p(statusOK), f64p(remainingPercent), C.double(threshold),
		&healthy, i32p(low), &nLow)
	if rc != 0 {
		return 0, nil, lastErr("klm_fleet_scan")
	}
	return int64(healthy), low[:int(nLow)], nil
}

// RingCreate allocates nSeries rows of `window` f32 samples in HBM.
func (k *Handle) RingCreate(nSeries int64, window int32) error {
	if rc := C.klm_ring_create(k.h, C.int64_t(nSeries), C.int32_t(window)); rc != 0 {
		return lastErr("klm_ring_create")
	}
	return nil
}

// RingIngest appends one sample to each listed series (UpdateUAVReport, manager.go:391-449).
func (k *Handle) RingIngest(seriesIDs []int64, values []float32) error {
	if rc := C.klm_ring_ingest(k.h, C.int64_t(len(seriesIDs)), i64p(seriesIDs), f32p(values)); rc != 0 {
		return lastErr("klm_ring_ingest")
	}
	return nil
}

// RingIngestTick appends one sample to every series (end of Collect, manager.go:286).
func (k *Handle) RingIngestTick(values []float32) error {
	if rc := C.klm_ring_ingest_tick(k.h, f32p(values)); rc != 0 {
		return lastErr("klm_ring_ingest_tick")
	}
	return nil
}

// WindowStats are min / max / mean / p95 of each series' window.
type WindowStats struct {
	Min, Max, P95 []float32
	Mean          []float64
}

// RingStats computes window statistics of series [s0, s1).
func (k *Handle) RingStats(s0, s1 int64) (*WindowStats, error) {
	n := int(s1 - s0)
	w := &WindowStats{make([]float32, n), make([]float32, n), make([]float32, n), make([]float64, n)}
	rc := C.klm_ring_stats(k.h, C.int64_t(s0), C.int64_t(s1), f32p(w.Min), f32p(w.Max), f64p(w.Mean), f32p(w.P95))
	if rc != 0 {
		return nil, lastErr("klm_ring_stats")
	}
	return w, nil
}

// HostFloat64s returns a slice of n float64 backed by page-locked C memory
// (klm_host_alloc). Arrays passed to the klm_* calls may be ordinary Go
// slices, but the driver stages those synchronously; page-locked arrays make
// the copies asynchronous, which lets klm_collect overlap its two branches and
// klm_schedule_batch pipeline a large batch over PCIe. Release with HostFree.
func HostFloat64s(n int) ([]float64, error) {
	var p unsafe.Pointer
	if rc := C.klm_host_alloc(C.size_t(n*8), &p); rc != 0 {
		return nil, lastErr("klm_host_alloc")
	}
	return unsafe.Slice((*float64)(p), n), nil
}

// HostFree releases a slice obtained from a Host* allocator.
func HostFree(p unsafe.Pointer) { C.klm_host_free(p) }

// CollectResult is one Manager.Collect cycle: cluster metrics, per-node rates and the
// window statistics of every series. Min / Max / P95 / Mean are views into one
// page-locked block owned by the Collector (valid until the next Collect).
type CollectResult struct {
	Cluster       ClusterOut
	CPURate       []float64
	MemoryRate    []float64
	Min, Max, P95 []float32
	Mean          []float64
}

// Collector holds the page-locked result block for klm_collect's packed output.
type Collector struct {
	k       *Handle
	nSeries int64
	block   unsafe.Pointer
}

// NewCollector prepares the result block for a ring of nSeries series (RingCreate first).
func (k *Handle) NewCollector(nSeries int64) (*Collector, error) {
	var p unsafe.Pointer
	if rc := C.klm_host_alloc(C.size_t(C.klm_collect_packed_bytes(C.int64_t(nSeries))), &p); rc != 0 {
		return nil, lastErr("klm_host_alloc")
	}
	return &Collector{k: k, nSeries: nSeries, block: p}, nil
}

// Close releases the result block.
func (c *Collector) Close() { C.klm_host_free(c.block); c.block = nil }

// Collect replaces the tail of Manager.Collect (manager.go:195-319): appends `tick` (one
// sample per series, or nil) to the ring and returns calculateClusterMetrics, the node
// rates and the window statistics from one call and one synchronisation.
func (c *Collector) Collect(tick []float32, cols *NodeCols, podRunning []uint8) (*CollectResult, error) {
	n := len(cols.CPUCapacity)
	res := &CollectResult{CPURate: make([]float64, n), MemoryRate: make([]float64, n)}
	var pin runtime.Pinner
	defer pin.Unpin()
	nc := cols.toC(&pin)
	var cl C.klm_cluster_out
	pinSlice(&pin, tick)
	pinSlice(&pin, podRunning)
	pinSlice(&pin, res.CPURate)
	pinSlice(&pin, res.MemoryRate)
	pin.Pin(&nc)
	pin.Pin(&cl)
	in := C.klm_collect_in{tick_values: f32p(tick), nodes: &nc, n_nodes: C.int64_t(n),
		pod_running: u8p(podRunning), n_pods: C.int64_t(len(podRunning))}
	out := C.klm_collect_out{cluster: &cl, cpu_rate: f64p(res.CPURate), memory_rate: f64p(res.MemoryRate),
		win_packed: c.block}
	if rc := C.klm_collect(c.k.h, &in, &out); rc != 0 {
		return nil, lastErr("klm_collect")
	}
	res.Cluster = clusterFromC(&cl)
	s := int(c.nSeries)
	c4 := uintptr((s*4 + 255) &^ 255)
	res.Min = unsafe.Slice((*float32)(c.block), s)
	res.Max = unsafe.Slice((*float32)(unsafe.Add(c.block, c4)), s)
	res.P95 = unsafe.Slice((*float32)(unsafe.Add(c.block, 2*c4)), s)
	res.Mean = unsafe.Slice((*float64)(unsafe.Add(c.block, 3*c4)), s)
	return res, nil
}
