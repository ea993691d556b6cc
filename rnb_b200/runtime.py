"""Worker loop of one pipeline step with the event-fenced hand-off of :mod:`rnb_b200.handoff`.

Same signature, queue protocol, termination rules, TimeCard stamps and log files as the
reference's ``runner()`` (``runner.py:5-270``), so ``benchmark.py`` can spawn it in place of the
reference loop (``scripts/run_rnb.py --runtime b200`` registers this module as ``runner``).  Only
the tensor-moving lines differ:

* inputs are pulled with ``handoff.fetch`` (``runner.py:107-117``): the copy waits, on the GPU, for
  the producer's ``ready`` event; the buffer is released to the producer at once and its next write
  waits for our ``free`` event, so neither side blocks a host thread on the other's GPU work;
* outputs are pushed with ``handoff.publish`` (``runner.py:140-173``): segments are cut with the
  reference's divmod rule, copied behind the step's own kernels, and announced downstream without
  the host-side ``stream.synchronize()`` of ``runner.py:127-128`` -- the consumer's copy is ordered
  by the event instead.  The host only synchronises where a host reads results: in the final step
  and for consumers without a GPU (the CPU aggregator);
* buffers are whatever ``control.SharedQueuesAndTensors`` built: with ``handoff.TensorEvent`` they
  carry a dtype (uint8 frames on the wire) and the events; with the reference's ``TensorEvent`` the
  helpers fall back to waiting for each copy, which is still correct.

Everything else -- queues, selectors, counters, barriers, logging -- is the reference's own code,
imported from the RnB checkout on ``sys.path``.
"""
NUM_EXIT_MARKERS = 10
NUM_SUMMARY_SKIPS = 10


def runner(input_queue, output_queues, queue_selector_path, print_summary,
           job_id, g_idx, group_idx, instance_idx,
           global_inference_counter, num_videos,
           termination_flag, step_idx,
           sta_bar, fin_bar,
           model_module_path, num_segments,
           shared_input_tensors, shared_output_tensors,
           **model_kwargs):
    import torch
    from queue import Full
    from tqdm import tqdm
    from rnb_logging import logname, TimeCardSummary, TimeCardList
    from control import TerminationFlag, Signal
    from utils.class_utils import load_class
    from rnb_b200 import handoff

    # RNB_B200_RUNTIME_WATCHDOG=<seconds>: dump every thread's Python stack to stderr if the worker
    # is still alive after that long (an exception in one RnB process leaves the others waiting at
    # the final barrier, benchmark.py:195-197; this shows who waits where)
    import os
    if os.environ.get('RNB_B200_RUNTIME_WATCHDOG'):
        import faulthandler
        faulthandler.enable()
        faulthandler.dump_traceback_later(float(os.environ['RNB_B200_RUNTIME_WATCHDOG']), exit=False)

    on_gpu = g_idx >= 0
    if on_gpu:
        torch.cuda.set_device(g_idx)
    device = torch.device('cuda:%d' % g_idx) if on_gpu else torch.device('cpu')
    stream = torch.cuda.Stream(device=g_idx) if on_gpu else None
    is_final_step = output_queues is None

    with torch.cuda.device(g_idx if on_gpu else None), torch.cuda.stream(stream), torch.no_grad():
        model = load_class(model_module_path)(device, **model_kwargs)
        if is_final_step:
            time_card_summary = TimeCardSummary()
        else:
            queue_selector = load_class(queue_selector_path)(len(output_queues))
        out_counter = 0

        placeholders = None
        if shared_input_tensors is not None:
            sample = list(shared_input_tensors.values())[0][0][0]
            placeholders = tuple(torch.zeros_like(t, device=device) for t in sample.tensors)

        sta_bar.wait()
        if print_summary:
            progress_bar = tqdm(total=num_videos)
            seen = 0

        while termination_flag.value == TerminationFlag.UNSET:
            tpl = input_queue.get()
            if tpl is None:
                break
            signal, non_tensor_inputs, time_card = tpl
            time_card.add_gpu(g_idx)
            time_card.record('runner%d_start' % step_idx)

            tensor_inputs = None
            if signal is not None:
                te = shared_input_tensors[signal.group_idx][signal.instance_idx][signal.tensor_idx]
                if te.event.is_set() and termination_flag.value != TerminationFlag.UNSET:
                    break
                if on_gpu:
                    tensor_inputs = [handoff.fetch(ph, te, index=i)
                                     for i, ph in enumerate(placeholders)]
                else:
                    # a step without a GPU (the aggregator of config/r2p1d-segment.json): wait for the
                    # producer's copy, then read the rows on the host
                    if getattr(te, 'ready', None) is not None:
                        te.ready.synchronize()
                    tensor_inputs = []
                    for i, (ph, shared) in enumerate(zip(placeholders, te.tensors)):
                        n = te.valid_batch_sizes[i]
                        ph[:n].copy_(shared[:n])
                        tensor_inputs.append(ph[:n])
                te.event.set()

            time_card.record('inference%d_start' % step_idx)
            tensor_outputs, non_tensor_outputs, time_card = \
                model(tensor_inputs, non_tensor_inputs, time_card)
            if time_card is None:
                continue
            if stream is not None and (is_final_step or shared_output_tensors is None):
                # a host consumer follows (log stamps of the last step, non-tensor results)
                stream.synchronize()
            time_card.record('inference%d_finish' % step_idx)

            if shared_output_tensors is not None:
                for seg_idx in range(num_segments):
                    te = shared_output_tensors[(out_counter + seg_idx) % len(shared_output_tensors)]
                    te.event.wait()
                    for i, out in enumerate(tensor_outputs):
                        lo, hi = handoff.segment_ranges(out.shape[0], num_segments)[seg_idx]
                        handoff.publish(te, out[lo:hi], index=i)
                    te.event.clear()

            if is_final_step:
                n_done = len(time_card.time_cards) if isinstance(time_card, TimeCardList) else 1
                with global_inference_counter.get_lock():
                    before = global_inference_counter.value
                    global_inference_counter.value = before + n_done
                    if print_summary and global_inference_counter.value > seen:
                        progress_bar.update(global_inference_counter.value - seen)
                        seen = global_inference_counter.value
                    if global_inference_counter.value >= num_videos:
                        if before < num_videos:
                            print('Finished processing %d videos' % num_videos)
                            termination_flag.value = TerminationFlag.TARGET_NUM_VIDEOS_REACHED
                        else:
                            break
                for tc in (time_card.time_cards if isinstance(time_card, TimeCardList)
                           else [time_card]):
                    time_card_summary.register(tc)
            else:
                queue = output_queues[queue_selector.select(tensor_outputs, non_tensor_outputs,
                                                            time_card)]
                try:
                    for seg_idx in range(num_segments):
                        tc = time_card.fork(seg_idx) if num_segments > 1 else time_card
                        sig = None
                        if shared_output_tensors is not None:
                            sig = Signal(group_idx, instance_idx, out_counter)
                            out_counter = (out_counter + 1) % len(shared_output_tensors)
                        queue.put_nowait((sig, non_tensor_outputs, tc))
                except Full:
                    print('[WARNING] Queue between runner step %d and %d is full. Aborting...'
                          % (step_idx, step_idx + 1))
                    termination_flag.value = TerminationFlag.FRAME_QUEUE_FULL
                    break

        if not is_final_step:
            try:
                for _ in range(NUM_EXIT_MARKERS):
                    for queue in output_queues:
                        queue.put_nowait(None)
            except Full:
                pass
        if shared_input_tensors is not None:
            for group_tensors in shared_input_tensors.values():
                for instance_tensors in group_tensors:
                    for te in instance_tensors:
                        te.event.set()
        if stream is not None:
            stream.synchronize()

    fin_bar.wait()
    if output_queues is not None:
        for queue in output_queues:
            queue.cancel_join_thread()
    if is_final_step:
        with open(logname(job_id, g_idx, group_idx, instance_idx), 'w') as f:
            time_card_summary.save_full_report(f)
        if print_summary:
            time_card_summary.print_summary(NUM_SUMMARY_SKIPS)
            progress_bar.close()
