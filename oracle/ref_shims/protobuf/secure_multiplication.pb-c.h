/* TEST INFRASTRUCTURE (oracle/_ref build only).
 * Hand-written equivalent of what protoc-c would generate from
 * /root/reference/src/protobuf/secure_multiplication.bitwidth_{32,64}.proto
 * (message msg { repeated uintW vector = 1 [packed=true]; required uintW value = 2; }),
 * restricted to the calls made at /root/reference/src/phase1.c:100-145,164,225-226,249-250.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

typedef struct ProtobufCBuffer ProtobufCBuffer;
struct ProtobufCBuffer {
	void (*append)(ProtobufCBuffer *buffer, size_t len, const uint8_t *data);
};
typedef struct { int unused; } ProtobufCMessage;
typedef struct { int unused; } ProtobufCAllocator;

#if BIT_WIDTH_32_P1
typedef uint32_t p1shim_word_t;
#else
typedef uint64_t p1shim_word_t;
#endif

typedef struct {
	ProtobufCMessage base;
	size_t n_vector;
	p1shim_word_t *vector;
	p1shim_word_t value;
} SecureMultiplication__Msg;

void secure_multiplication__msg__init(SecureMultiplication__Msg *m);
size_t secure_multiplication__msg__get_packed_size(const SecureMultiplication__Msg *m);
size_t secure_multiplication__msg__pack_to_buffer(const SecureMultiplication__Msg *m, ProtobufCBuffer *buffer);
SecureMultiplication__Msg *secure_multiplication__msg__unpack(ProtobufCAllocator *a, size_t len, const uint8_t *data);
void secure_multiplication__msg__free_unpacked(SecureMultiplication__Msg *m, ProtobufCAllocator *a);
