// Package drift — canon.go (UNCOMPILED, see go/README.md).
//
// Canonical encoding of a Go value such that
//
//	bytes.Equal(Canon(a), Canon(b))  <=>  reflect.DeepEqual(a, b)
//
// for the value shapes the controller compares (controller.go:539,600,725,795): structs,
// pointers, slices, maps, strings, []byte, bools, integers.  Same format as
// oracle/go_values.cpp and csrc/host/go_canon.cpp:
//
//	node    = tag byte, content
//	lengths = uint32 little endian; integers = width byte + int64 little endian
//	nil map / slice / []byte / pointer / interface have their own tags (nil != empty)
//	map     = entries sorted by the canonical bytes of the key
//	float   = IEEE bits, -0 folded into +0; NaN is an error (NaN != NaN)
package drift

import (
	"bytes"
	"encoding/binary"
	"errors"
	"math"
	"reflect"
	"sort"
)

const (
	tBool byte = iota + 1
	tInt
	tString
	tBytes
	tNilBytes
	tSlice
	tNilSlice
	tMap
	tNilMap
	tPtr
	tNilPtr
	tStruct
	tFloat
	tIface
	tNilIface
)

var errNaN = errors.New("drift: NaN cannot be canonicalised (it is not DeepEqual to itself)")

// Canon appends the canonical encoding of v to dst.
func Canon(dst []byte, v reflect.Value) ([]byte, error) {
	switch v.Kind() {
	case reflect.Bool:
		b := byte(0)
		if v.Bool() {
			b = 1
		}
		return append(dst, tBool, b), nil
	case reflect.Int, reflect.Int8, reflect.Int16, reflect.Int32, reflect.Int64:
		dst = append(dst, tInt, byte(v.Type().Bits()))
		return binary.LittleEndian.AppendUint64(dst, uint64(v.Int())), nil
	case reflect.Uint, reflect.Uint8, reflect.Uint16, reflect.Uint32, reflect.Uint64, reflect.Uintptr:
		dst = append(dst, tInt, byte(v.Type().Bits())|0x80)
		return binary.LittleEndian.AppendUint64(dst, v.Uint()), nil
	case reflect.Float32, reflect.Float64:
		f := v.Float()
		if math.IsNaN(f) {
			return dst, errNaN
		}
		if f == 0 {
			f = 0 // folds -0 into +0: they are ==, hence DeepEqual
		}
		dst = append(dst, tFloat)
		return binary.LittleEndian.AppendUint64(dst, math.Float64bits(f)), nil
	case reflect.String:
		dst = append(dst, tString)
		dst = binary.LittleEndian.AppendUint32(dst, uint32(v.Len()))
		return append(dst, v.String()...), nil
	case reflect.Slice:
		if v.Type().Elem().Kind() == reflect.Uint8 {
			if v.IsNil() {
				return append(dst, tNilBytes), nil
			}
			dst = append(dst, tBytes)
			dst = binary.LittleEndian.AppendUint32(dst, uint32(v.Len()))
			return append(dst, v.Bytes()...), nil
		}
		if v.IsNil() {
			return append(dst, tNilSlice), nil
		}
		fallthrough
	case reflect.Array:
		dst = append(dst, tSlice)
		dst = binary.LittleEndian.AppendUint32(dst, uint32(v.Len()))
		var err error
		for i := 0; i < v.Len(); i++ {
			if dst, err = Canon(dst, v.Index(i)); err != nil {
				return dst, err
			}
		}
		return dst, nil
	case reflect.Map:
		if v.IsNil() {
			return append(dst, tNilMap), nil
		}
		type entry struct{ k, v []byte }
		entries := make([]entry, 0, v.Len())
		iter := v.MapRange()
		for iter.Next() {
			k, err := Canon(nil, iter.Key())
			if err != nil {
				return dst, err
			}
			val, err := Canon(nil, iter.Value())
			if err != nil {
				return dst, err
			}
			entries = append(entries, entry{k, val})
		}
		sort.Slice(entries, func(i, j int) bool { return bytes.Compare(entries[i].k, entries[j].k) < 0 })
		dst = append(dst, tMap)
		dst = binary.LittleEndian.AppendUint32(dst, uint32(len(entries)))
		for _, e := range entries {
			dst = append(dst, e.k...)
			dst = append(dst, e.v...)
		}
		return dst, nil
	case reflect.Pointer:
		if v.IsNil() {
			return append(dst, tNilPtr), nil
		}
		return Canon(append(dst, tPtr), v.Elem())
	case reflect.Interface:
		if v.IsNil() {
			return append(dst, tNilIface), nil
		}
		name := v.Elem().Type().String()
		dst = append(dst, tIface)
		dst = binary.LittleEndian.AppendUint32(dst, uint32(len(name)))
		dst = append(dst, name...)
		return Canon(dst, v.Elem())
	case reflect.Struct:
		dst = append(dst, tStruct)
		dst = binary.LittleEndian.AppendUint32(dst, uint32(v.NumField()))
		var err error
		for i := 0; i < v.NumField(); i++ { // unexported fields included, as DeepEqual does
			if dst, err = Canon(dst, v.Field(i)); err != nil {
				return dst, err
			}
		}
		return dst, nil
	}
	return dst, errors.New("drift: unsupported kind " + v.Kind().String())
}
