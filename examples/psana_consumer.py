"""Consumer example, the counterpart of the reference's examples/psana_consumer.py (ref: :1-55).

Launch the producer ranks (one per GPU; ranks come from mpirun or from RANK/WORLD_SIZE):
    mpirun -n 4 psana-ray-producer --exp mfxl1038923 --run 58 --detector_name epix10k2M --queue_size 400
Launch a consumer on the same box:
    python psana_consumer.py 1

Differences from the reference example: no Ray head to start (the producer's rank 0 starts the queue
server); the queue name/namespace are passed explicitly because the reference's defaults do not match
each other (producer: 'my'/'default', DataReader: 'shared_queue'/'my' -- ref: producer.py:26-27 vs
data_reader.py:5); an item is unpacked into its FOUR fields (the reference unpacks three and therefore
never processes a frame, ref: examples/psana_consumer.py:35 vs producer.py:101); `data` is a CUDA tensor.
"""
import signal
import sys
import time

from psana_ray.data_reader import DataReader, DataReaderError


def signal_handler(sig, frame):
    print("Ctrl+C pressed. Shutting down...")
    sys.exit(0)


def consume_data(consumer_id, queue_name="my", ray_namespace="default"):
    with DataReader(queue_name=queue_name, ray_namespace=ray_namespace) as reader:
        while True:
            try:
                result = reader.read()
                if result is not None:
                    rank, idx, data, photon_energy = result
                    print(f"Consumer {consumer_id} processed: rank={rank} | idx={idx} | shape={tuple(data.shape)} | device={data.device}")
                elif reader.end_of_stream:
                    print(f"Consumer {consumer_id}: end of stream")
                    break
                else:
                    print(f"Consumer {consumer_id} waiting for data...")
                    time.sleep(1)
            except DataReaderError as e:
                print(f"DataReader error: {e}")
                print("Queue actor is dead. Exiting...")
                break


if __name__ == "__main__":
    signal.signal(signal.SIGINT, signal_handler)
    consume_data(sys.argv[1] if len(sys.argv) > 1 else 1)
