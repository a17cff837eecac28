"""Consumer example -- what the reference's examples/psana_consumer.py does (ref: :28-47), written
against this package's additions.

    # producers, one rank per GPU (ranks from mpirun or RANK/WORLD_SIZE); rank 0 starts the queue server
    mpirun -n 4 psana-ray-producer --exp mfxl1038923 --run 58 --detector_name epix10k2M --queue_size 400
    # any number of consumers on the same box
    python examples/psana_consumer.py --id 1

Notes for users of the reference example:
  * there is no Ray head to start;
  * pass the producer's queue name and namespace ('my' / 'default'): the reference's own defaults
    do not agree with each other (ref: producer.py:26-27 vs data_reader.py:5);
  * an item has FOUR fields -- the reference example unpacks three and so never gets to a frame
    (ref: examples/psana_consumer.py:35 vs producer.py:101);
  * `data` is a float32 CUDA tensor mapped straight from the producer's GPU memory;
  * iterating a DataReader ends at the producer's end-of-stream sentinel, so no sleep-and-poll loop.
"""
import argparse

from psana_ray.data_reader import DataReader, DataReaderError


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--id", default="1")
    ap.add_argument("--queue_name", default="my")
    ap.add_argument("--ray_namespace", default="default")
    ap.add_argument("--batches", action="store_true", help="take whole GPU batches instead of single events")
    a = ap.parse_args()
    nevents = 0
    try:
        with DataReader(queue_name=a.queue_name, ray_namespace=a.ray_namespace) as reader:
            if a.batches:
                while not reader.end_of_stream:
                    batch = reader.read_batch()
                    if batch is None:
                        continue
                    rank, idx0, data, energies = batch
                    nevents += data.shape[0]
                    print(f"consumer {a.id}: rank {rank} events {idx0}..{idx0 + data.shape[0] - 1} {tuple(data.shape)} on {data.device}")
            else:
                for rank, idx, data, photon_energy in reader:
                    nevents += 1
                    print(f"consumer {a.id}: rank {rank} event {idx} {tuple(data.shape)} E={photon_energy} sum={float(data.sum()):.1f}")
    except DataReaderError as err:
        print(f"consumer {a.id}: the queue is gone ({err})")
    except KeyboardInterrupt:
        pass
    print(f"consumer {a.id}: {nevents} events")


if __name__ == "__main__":
    main()
